#!/bin/bash
# Round 2, GPU call 2: full GPU suite on the rebuilt library (sliced-ELL generic CSR, remap entry, merged reductions), the generic SpMV probe with both
# kernel families, the bench lines of the other BASELINE configs, and `--set full` captures of the generic SpMV kernels.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -x -q --durations=10 > $O/r02_pytest_gpu_2.log 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_2.log
timeout 600 python scripts/spmv_probe.py > $O/r02_spmv_probe_sell.log 2>&1; echo "probe sell rc=$?"
UFE_B200_CSR_KERNEL=vector timeout 600 python scripts/spmv_probe.py > $O/r02_spmv_probe_vector.log 2>&1; echo "probe vector rc=$?"
for cfg in mismip8km mismipplus_ssa; do
  timeout 900 python bench.py --config $cfg --steps 10 --warmup 3 --no-cpu-baseline --verbose > $O/r02_bench_$cfg.log 2> $O/r02_bench_$cfg.err; echo "bench $cfg rc=$?"
done
for nv in 1000000 4000000 16000000; do
  timeout 900 python bench.py --config sweep --nv $nv --steps 3 --warmup 3 --verbose > $O/r02_bench_sweep_${nv}.log 2> $O/r02_bench_sweep_${nv}.err; echo "sweep $nv rc=$?"
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_spmv_fused|k_sell_multicol' --launch-skip 16 -c 4 -o $O/r02_generic_spmv_4M -f \
  python scripts/spmv_probe.py > $O/r02_ncu_spmv.log 2>&1; echo "ncu spmv rc=$?"
ls -la $O
