#!/usr/bin/env python
"""BASELINE config 1 on this machine's CPUs: 1 ps + 2 workers on localhost, async and sync, gloo vs the native tcp
parameter server. Prints global steps per second of the cluster (wall clock of the slowest worker)."""
import json
import os
import re
import socket
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def run(transport, sync, steps=600):
    ports = free_ports(3)
    d = tempfile.mkdtemp()
    base = [sys.executable, "-m", "tensorflow_train_distributed_b200.train", f"--ps_hosts=127.0.0.1:{ports[0]}",
            f"--worker_hosts=127.0.0.1:{ports[1]},127.0.0.1:{ports[2]}", "--device=cpu", f"--transport={transport}",
            "--dataset_size=8192", f"--training_steps={steps}", "--save_checkpoint_secs=none", "--write_events=false",
            f"--checkpoint_dir={d}", "--log_every_n_steps=100"] + (["--sync_replicas"] if sync else [])
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2", MKL_NUM_THREADS="2")
    procs = [subprocess.Popen(base + ["--job_name=ps", "--task_id=0"], env=env, cwd=ROOT, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True)]
    for i in range(2):
        procs.append(subprocess.Popen(base + ["--job_name=worker", f"--task_id={i}"], env=env, cwd=ROOT,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=600)[0] for p in procs[1:]]
    procs[0].communicate(timeout=60)
    rates = []
    for i in range(2):
        rows = [json.loads(l) for l in open(os.path.join(d, f"metrics-{i}.jsonl"))]
        r = [x["steps_per_sec"] for x in rows if x["steps_per_sec"]]
        rates.append(sum(r) / max(len(r), 1))
    gs = [int(re.search(r"global_step: (\d+)", o).group(1)) for o in outs]
    return {"transport": transport, "mode": "sync" if sync else "async", "local_steps_per_sec_per_worker": [round(x, 1) for x in rates],
            "final_global_step": gs}


if __name__ == "__main__":
    for transport in ("gloo", "tcp"):
        for sync in (False, True):
            t0 = time.time()
            print("CPUCLUSTER " + json.dumps(dict(run(transport, sync), wall_s=round(time.time() - t0, 1))), flush=True)
