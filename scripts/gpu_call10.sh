#!/bin/bash
# Round 2, final one-GPU call: full GPU suite on the final library, the default bench line (with the CPU arm), its ncu launch list.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -x -q --durations=5 > $O/r02_pytest_gpu_final.log 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_final.log
timeout 1500 python bench.py > $O/r02_bench_final.log 2> $O/r02_bench_final.err; rc=$?; echo "bench rc=$rc"
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
if [ $rc -eq 0 ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 30000 --csv --log-file $O/r02_launches_raw_final.csv $B > $O/r02_ncu_list_final.log 2>&1
  echo "ncu list rc=$?"
fi
ls -la $O
