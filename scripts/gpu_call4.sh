#!/bin/bash
# Round 2, GPU call 4 (one GPU): full GPU suite, multigrid hierarchy probes (where to stop coarsening, how many sweeps on the coarsest level) on three
# meshes, generic multi-column SpMV with and without the column-interleaved right-hand sides, configs[0] with the Jacobi preconditioner.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -x -q --durations=8 > $O/r02_pytest_gpu_4.log 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_4.log
D="1,3,1.5,0.8,0"
S="$D;$D,stop_nodes=2000,coarse_sweeps=12;$D,stop_nodes=2000,coarse_sweeps=16;$D,stop_nodes=2000,coarse_sweeps=24;$D,stop_nodes=2000,coarse_sweeps=32;1,2,1.5,0.8,0,stop_nodes=2000,coarse_sweeps=16;1,2,1.5,0.8,0,stop_nodes=2000,coarse_sweeps=32;$D,stop_nodes=12000,coarse_sweeps=32;$D,stop_nodes=12000,coarse_sweeps=64;$D,stop_nodes=300,coarse_sweeps=16"
timeout 900 python scripts/mg_probe.py 4000000 3 "$S" > $O/r02_mg_probe_stop_4M.log 2>&1; echo "mg probe 4M rc=$?"
timeout 300 python scripts/mg_probe.py 72000 4 "jacobi;$S;$D,fuse=1;$D,tail_max=256;$D,tail_max=2048;1,1,1.5,0.8,0;1,0,1.5,0.8,0" > $O/r02_mg_probe_stop_72k.log 2>&1; echo "mg probe 72k rc=$?"
PROBE_WORKLOAD=mismip_plus timeout 600 python scripts/mg_probe.py 500000 3 "$S" > $O/r02_mg_probe_stop_mplus.log 2>&1; echo "mg probe mismip+ rc=$?"
timeout 600 python scripts/spmv_probe.py > $O/r02_spmv_probe_sell_v2.log 2>&1; echo "spmv probe rc=$?"
timeout 600 python bench.py --config mismip8km --pc jacobi --steps 10 --warmup 3 --no-cpu-baseline --verbose > $O/r02_bench_mismip8km_jacobi.log 2> $O/r02_bench_mismip8km_jacobi.err; echo "bench 8km jacobi rc=$?"
ls -la $O
