#!/bin/bash
# Round 2, GPU call 8 (one GPU): the DSMEM variant of the coarsest-level cluster kernel -- tests, probe at 4 M vertices, default bench line.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_multigrid.py tests/test_zz_no_alloc.py -m gpu -q > $O/r02_pytest_gpu_8.log 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_8.log
D="1,3,1.5,0.8,0"; T="stop_nodes=2000,coarsest_cluster=16"
S="$D,$T,coarse_sweeps=32,coarsest_dsmem=0;$D,$T,coarse_sweeps=32;$D,$T,coarse_sweeps=48;$D,$T,coarse_sweeps=64;$D,$T,coarse_sweeps=96;$D,stop_nodes=2000,coarsest_cluster=8,coarse_sweeps=48"
timeout 900 python scripts/mg_probe.py 4000000 3 "$S" > $O/r02_mg_probe_coarsest_dsmem_4M.log 2>&1; echo "mg probe rc=$?"
timeout 900 python bench.py --no-cpu-baseline > $O/r02_bench_5_dsmem.log 2> $O/r02_bench_5_dsmem.err; echo "bench rc=$?"
timeout 900 python bench.py --no-cpu-baseline --mg-coarse-sweeps 48 > $O/r02_bench_5_dsmem_cs48.log 2> $O/r02_bench_5_dsmem_cs48.err; echo "bench cs48 rc=$?"
ls -la $O
