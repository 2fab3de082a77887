#!/bin/bash
# Round 2, GPU call 7 (one GPU): the coarsest-level cluster kernel -- its test, the probe at 4 M vertices, and the bench line with it.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_multigrid.py tests/test_gpu_linear_algebra.py tests/test_zz_no_alloc.py -m gpu -q > $O/r02_pytest_gpu_7.log 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_7.log
D="1,3,1.5,0.8,0"; T="stop_nodes=2000"
S="$D,$T,coarse_sweeps=24;$D,$T,coarse_sweeps=24,coarsest_cluster=16;$D,$T,coarse_sweeps=24,coarsest_cluster=8;$D,$T,coarse_sweeps=32,coarsest_cluster=16;$D,$T,coarse_sweeps=48,coarsest_cluster=16;$D,$T,coarse_sweeps=64,coarsest_cluster=16;$D,$T,coarse_sweeps=96,coarsest_cluster=16"
timeout 900 python scripts/mg_probe.py 4000000 3 "$S" > $O/r02_mg_probe_coarsest_cluster_4M.log 2>&1; echo "mg probe rc=$?"
timeout 900 python bench.py --mg-coarsest-cluster 16 --mg-coarse-sweeps 32 --no-cpu-baseline > $O/r02_bench_4_cluster16_cs32.log 2> $O/r02_bench_4_cluster16_cs32.err; echo "bench rc=$?"
ls -la $O
