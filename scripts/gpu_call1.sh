#!/bin/bash
# Round 2, GPU call 1: the no-alloc test alone and after the suite, the full GPU suite, the default bench line, the ncu launch list of one timed
# step and one `--set full` capture of the first launch of each kernel of the step.  Every ncu pass only after the same command exited 0 without ncu.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > $O/r02_box.txt 2>&1
timeout 900 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/r02_pytest_no_alloc_alone.log 2>&1; echo "rc=$?" >> $O/r02_pytest_no_alloc_alone.log
timeout 3000 python -m pytest tests -m gpu -x -q --durations=15 > $O/r02_pytest_gpu_1.log 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_1.log
timeout 1200 python bench.py > $O/r02_bench_1.log 2> $O/r02_bench_1.err; rc=$?; echo "bench rc=$rc"
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $B > $O/r02_bench_1step.log 2> $O/r02_bench_1step.err; rc1=$?; echo "bench 1 step rc=$rc1"
if [ $rc1 -eq 0 ]; then
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 30000 --csv --log-file $O/r02_launches_raw.csv $B > $O/r02_ncu_list.log 2>&1
  echo "ncu list rc=$?"
  K='regex:k_tri_assemble|k_vertex_diva|k_epilogue|k_th_flux|k_th_semiimplicit|k_mg_galerkin|k_ks_update|k_ksmg_|k_spmv_blk|k_spmv_fused|k_mg_sweep|k_mg_smooth0|k_mg_restrict|k_mg_prolong|k_geo_masks|k_geo_frac|k_geo_hi_eff|k_pc_'
  timeout 1500 ncu --set full --clock-control none --import-source on --profile-from-start off --kernel-id "::$K:1" -o $O/r02_step_kernels_4M -f $B > $O/r02_ncu_full.log 2>&1
  echo "ncu full rc=$?"
fi
ls -la $O
