#!/bin/bash
# Round 2, GPU call 11 (one GPU): smoke() of __graft_entry__, and the coarsest-level cluster kernel with fewer CTAs / threads.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/r02_smoke.log
D="1,3,1.5,0.8,0"; T="stop_nodes=2000,coarse_sweeps=32"
S="$D,$T,coarsest_cluster=16;$D,$T,coarsest_cluster=16,coarsest_threads=96;$D,$T,coarsest_cluster=16,coarsest_threads=128;$D,$T,coarsest_cluster=4,coarsest_threads=256;$D,$T,coarsest_cluster=4,coarsest_dsmem=1;$D,$T,coarsest_cluster=8,coarsest_dsmem=1,coarsest_threads=128;$D,$T,coarsest_cluster=16,coarsest_dsmem=1,coarsest_threads=96"
timeout 600 python scripts/mg_probe.py 4000000 3 "$S" > $O/r02_mg_probe_coarsest_threads_4M.log 2>&1; echo "mg probe rc=$?"
