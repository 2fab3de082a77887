#!/bin/bash
# All multi-GPU correctness checks in one run:  scripts/multi_gpu_check.sh N [logfile]   (N = 2, 4 or 8 GPUs of one box)
#   1. multi_gpu_parity.py   DIVA solve + 2 predictor-corrector steps, Jacobi and multigrid: N ranks against one GPU (<= 1e-6) and against the CPU oracle
#   2. multi_gpu_copy_bc.py  copy boundary conditions whose source rows live on another rank
#   3. stage_parity_bench.py geometry / mask / sub-grid / secondary-velocity stages: integer masks identical, fields <= 1e-12
# Exit code 0 only if all three pass.
cd "$(dirname "$0")/.." || exit 1
N=${1:-2}; LOG=${2:-gpurun_out/r02_multi_gpu_check_${N}gpu.log}
mkdir -p "$(dirname "$LOG")"; : > "$LOG"
rc=0; port=29540
for s in multi_gpu_parity.py multi_gpu_copy_bc.py stage_parity_bench.py; do
  echo "== $s on $N GPUs" >> "$LOG"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port $port scripts/$s >> "$LOG" 2>&1
  r=$?; echo "== $s exit code $r" >> "$LOG"; [ $r -ne 0 ] && rc=1
  port=$((port + 1))
done
echo "== all checks: $([ $rc -eq 0 ] && echo PASSED || echo FAILED)" >> "$LOG"
exit $rc
