#!/bin/bash
# 2-GPU regression: GPU tests, sync parity (NVLS forced and off), K1 timing. Usage: gpurun --gpus 2 -- bash tools/two_gpu_check.sh
P=29700
run() { P=$((P + 1)); timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $P "$@"; }
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for nv in 1 auto; do
  for m in mlp lenet; do
    TTD_NVLS=$nv CHECK_MODEL=$m run tools/mgpu_check.py 2>&1 | grep "MGPU\|Error\|error" | head -3
  done
  TTD_NVLS=$nv CHECK_PULL_FUSED=1 run tools/mgpu_check.py 2>&1 | grep "MGPU\|Error\|error" | head -3
  TTD_NVLS=$nv run tools/k1_bench.py 2>&1 | grep "K1BENCH\|Error" | head -3
done
