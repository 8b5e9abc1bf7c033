#!/bin/bash
# Multi-GPU validation + scaling sweep on one box: `gpurun --gpus 8 -- bash tools/scale_run.sh 8`.
# Every step runs under its own timeout; results land in gpurun_out/scale/.
NMAX=${1:-8}
OUT=gpurun_out/scale
mkdir -p $OUT
PORT=29600
tr() {  # tr <nproc> <script> [args...]
  local n=$1; shift
  PORT=$((PORT + 1))
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
    --master-port $PORT "$@"
}
nvidia-smi topo -m > $OUT/topo.txt 2>&1
# 1. correctness at NMAX GPUs: sync parity against the single-GPU global batch, async atomics, LeNet, wide
for mode in sync async; do
  [ -n "$SKIP_CHECKS" ] && break
  for model in ${CHECK_MODELS:-mlp}; do
    CHECK_MODE=$mode CHECK_MODEL=$model tr $NMAX tools/mgpu_check.py > $OUT/check_${mode}_${model}_n$NMAX.log 2>&1
    echo "check $mode $model rc=$?"; tail -2 $OUT/check_${mode}_${model}_n$NMAX.log
  done
done
# 2. K1 (reduce + update + broadcast) with and without NVLS
for model in ${K1_MODELS-mlp wide_mlp}; do
  for nvls in auto 0; do
    [ $model = wide_mlp ] && [ $nvls = 0 ] && [ -z "$FULL" ] && continue
    TTD_NVLS=$nvls K1_MODEL=$model tr $NMAX tools/k1_bench.py > $OUT/k1_${model}_nvls${nvls}_n$NMAX.log 2>&1
    echo "k1 $model nvls=$nvls rc=$?"; tail -1 $OUT/k1_${model}_nvls${nvls}_n$NMAX.log
  done
done
# 3. scaling sweep of the headline config, ours and the torch DDP arm
for n in ${NS:-1 2 4 8}; do
  [ $n -gt $NMAX ] && break
  cfgs="mnist_sync mnist_async wide_sync lenet_sync"
  [ $n -ne $NMAX ] && [ -z "$FULL" ] && cfgs="mnist_sync"
  [ -n "$CFGS" ] && cfgs="$CFGS"
  for cfg in $cfgs; do
    steps=2000; [ $cfg = wide_sync ] && steps=200; [ $cfg = lenet_sync ] && steps=500
    if [ $n -eq 1 ]; then
      timeout 240 python bench.py --config $cfg --gpus 1 --steps $steps --warmup 20 > $OUT/bench_${cfg}_n$n.json 2> $OUT/bench_${cfg}_n$n.err
    else
      tr $n bench.py --config $cfg --gpus $n --steps $steps --warmup 20 > $OUT/bench_${cfg}_n$n.json 2> $OUT/bench_${cfg}_n$n.err
    fi
    echo "bench $cfg n=$n rc=$?"; tail -c 400 $OUT/bench_${cfg}_n$n.json | head -c 0
    python - <<EOF
import json
try:
    d = json.loads(open("$OUT/bench_${cfg}_n$n.json").read().strip().splitlines()[-1])
    print("  ", d["value"], "ex/s", d["ms_per_step"] * 1e3, "us/step  e2e", d["e2e"]["value"])
except Exception as e:
    print("   no result:", e)
EOF
  done
  if [ $n -eq $NMAX ] && [ $n -gt 1 ]; then  # K3 off: W1^T broadcast by K1 instead of pulled by the first GEMM
    tr $n bench.py --config mnist_sync --gpus $n --steps 2000 --warmup 20 --no-pull-fused > $OUT/bench_mnist_sync_nopull_n$n.json 2> $OUT/bench_mnist_sync_nopull_n$n.err
    python -c "import json; d=json.loads(open('$OUT/bench_mnist_sync_nopull_n$n.json').read().strip().splitlines()[-1]); print('no-pull-fused n=$n', d['ms_per_step']*1e3, 'us/step')"
  fi
  [ -n "$SKIP_DDP" ] && continue
  [ $n -ne $NMAX ] && [ -z "$FULL" ] && continue
  if [ $n -eq 1 ]; then
    timeout 240 python bench.py --impl ddp --gpus 1 --steps 500 --warmup 20 > $OUT/ddp_n$n.json 2> $OUT/ddp_n$n.err
  else
    tr $n bench.py --impl ddp --gpus $n --steps 500 --warmup 20 > $OUT/ddp_n$n.json 2> $OUT/ddp_n$n.err
  fi
  echo "ddp n=$n rc=$?"; tail -c 300 $OUT/ddp_n$n.json
done
echo done
