#!/bin/bash
# Throughput of the user-facing CLI loop (session + hooks + reader), three ways. 1 GPU.
run() {
  local tag=$1; shift
  local d=/tmp/cli_tp_$tag
  rm -rf $d
  python -m tensorflow_train_distributed_b200.train --job_name=worker --task_id=0 --worker_hosts=127.0.0.1:2222 \
    --ps_hosts=127.0.0.1:2221 --sync_replicas --training_steps=20000 --save_checkpoint_secs=none \
    --checkpoint_dir=$d --log_every_n_steps=5000 "$@" 2>&1 | grep -E "total step|Error|error" | sed "s/^/[$tag] /"
  python - <<PY
import json
rows = [json.loads(l) for l in open("$d/metrics-0.jsonl")]
r = [x["steps_per_sec"] for x in rows if x["steps_per_sec"]]
print("[$tag] steps/s per logging period:", [round(v) for v in r], "-> us/step", round(1e6 / (sum(r[1:]) / max(len(r[1:]), 1)), 1) if len(r) > 1 else None)
PY
}
run exact
run lag8 --fetch_lag=8
run lag8_prefetch4 --fetch_lag=8 --prefetch=4
run prefetch4 --prefetch=4
