#!/bin/bash
# Round 2, last one-GPU call: full GPU suite on the final library and tests.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q --durations=3 > $O/r02_pytest_gpu_final2.log 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_final2.log
tail -3 $O/r02_pytest_gpu_final2.log
