#!/bin/bash
# Round 2, GPU call 5 (one GPU): GPU suite, the coarsest-level sweeps as one single-CTA / cluster kernel per visit, the default bench line with the
# hierarchy options bench.py now passes for this mesh family, and its ncu launch list.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -x -q --durations=5 > $O/r02_pytest_gpu_5.log 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_5.log
D="1,3,1.5,0.8,0"; T="stop_nodes=2000,tail_max=2000"
S="$D,stop_nodes=2000,coarse_sweeps=24;$D,$T,coarse_sweeps=24;$D,$T,coarse_sweeps=24,tail_cluster=8;$D,$T,coarse_sweeps=24,tail_cluster=16;$D,$T,coarse_sweeps=32,tail_cluster=8;$D,$T,coarse_sweeps=48,tail_cluster=16;$D,$T,coarse_sweeps=64,tail_cluster=16;$D,stop_nodes=2000,coarse_sweeps=20;$D,stop_nodes=2000,coarse_sweeps=28"
timeout 900 python scripts/mg_probe.py 4000000 3 "$S" > $O/r02_mg_probe_cluster_4M.log 2>&1; echo "mg probe rc=$?"
timeout 1500 python bench.py > $O/r02_bench_3.log 2> $O/r02_bench_3.err; rc=$?; echo "bench rc=$rc"
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 30000 --csv --log-file $O/r02_launches_raw_v2.csv $B > $O/r02_ncu_list_v2.log 2>&1
  echo "ncu list rc=$?"
fi
ls -la $O
