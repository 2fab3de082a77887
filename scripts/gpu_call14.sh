#!/bin/bash
# Round 2, GPU call 14 (one GPU): how few Chebyshev steps on the coarsest level are enough; bench line with 8 steps.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
D="1,3,1.5,0.8,0"; T="stop_nodes=2000,coarsest_cluster=16,coarsest_cheb=1"
S="$D,$T,coarse_sweeps=8;$D,$T,coarse_sweeps=6;$D,$T,coarse_sweeps=4;$D,$T,coarse_sweeps=3;$D,$T,coarse_sweeps=6,cheb_ratio=10;$D,$T,coarse_sweeps=8,cheb_ratio=15;$D,$T,coarse_sweeps=10,cheb_ratio=60;1,2,1.5,0.8,0,$T,coarse_sweeps=8;$D,stop_nodes=12000,coarsest_cluster=16,coarsest_cheb=1,coarse_sweeps=16,cheb_ratio=100"
timeout 400 python scripts/mg_probe.py 4000000 3 "$S" > $O/r02_mg_probe_cheb2_4M.log 2>&1; echo "mg probe rc=$?"
timeout 600 python bench.py --mg-coarsest-cheb 1 --mg-coarse-sweeps 8 --no-cpu-baseline > $O/r02_bench_6_cheb8.log 2> $O/r02_bench_6_cheb8.err; echo "bench rc=$?"
