#!/bin/bash
# Round 2, multi-GPU call: scripts/gpu_call6.sh N  (run under `gpurun --gpus N`).  All multi-GPU correctness checks (against one GPU and against the
# CPU oracle), then the default bench line on N GPUs.
cd "$(dirname "$0")/.." || exit 1
N=${1:-2}; O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/r02_topo_${N}gpu.txt 2>&1
PARITY_ONLY=${PARITY_ONLY:-0}
if [ "$PARITY_ONLY" = "1" ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29540 scripts/multi_gpu_parity.py > $O/r02_multi_gpu_parity_${N}gpu.log 2>&1
  echo "parity rc=$?"
else
  bash scripts/multi_gpu_check.sh $N $O/r02_multi_gpu_check_${N}gpu.log; echo "multi-GPU check rc=$?"
fi
[ "${SKIP_BENCH:-0}" = "1" ] || timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29550 bench.py --gpus $N --steps 5 --warmup 3 \
  > $O/r02_bench_${N}gpu.log 2> $O/r02_bench_${N}gpu.err; echo "bench rc=$?"
tail -c 600 $O/r02_multi_gpu_*_${N}gpu.log
ls -la $O
