#!/bin/bash
# Round 2, GPU call 13 (one GPU): Chebyshev steps on the coarsest level -- test and probe at 4 M vertices.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_multigrid.py -m gpu -q -k "chebyshev or cluster_kernel" > $O/r02_pytest_gpu_13.log 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_13.log
D="1,3,1.5,0.8,0"; T="stop_nodes=2000,coarsest_cluster=16"
S="$D,$T,coarse_sweeps=32;$D,$T,coarsest_cheb=1,coarse_sweeps=8;$D,$T,coarsest_cheb=1,coarse_sweeps=12;$D,$T,coarsest_cheb=1,coarse_sweeps=16;$D,$T,coarsest_cheb=1,coarse_sweeps=24;$D,$T,coarsest_cheb=1,coarse_sweeps=12,cheb_ratio=10;$D,$T,coarsest_cheb=1,coarse_sweeps=12,cheb_ratio=100;$D,$T,coarsest_cheb=1,coarse_sweeps=16,cheb_ratio=100;$D,$T,coarsest_cheb=1,coarse_sweeps=24,cheb_ratio=300"
timeout 400 python scripts/mg_probe.py 4000000 3 "$S" > $O/r02_mg_probe_cheb_4M.log 2>&1; echo "mg probe rc=$?"
