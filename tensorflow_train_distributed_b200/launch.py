"""Spawn a whole cluster on this machine — what the reference's operator does by hand with one terminal per
``python distribute_training.py --job_name=... --task_id=...`` (README.md, flag defaults at :24-36).

    python -m tensorflow_train_distributed_b200.launch --ps_hosts=127.0.0.1:2221 \
        --worker_hosts=127.0.0.1:2222,127.0.0.1:2223 [any training flag ...]

Every ps and worker task of the cluster becomes one subprocess of ``...train`` with its ``--job_name`` /
``--task_id``; worker output is prefixed with its task. The exit code is the worst worker exit code; ps tasks
that outlive the workers (the reference's never exit, :181) are terminated once the workers are done.
"""
from __future__ import annotations

import subprocess
import sys
import threading
from typing import List, Optional, Sequence

from .config import parse_flags


def _pump(prefix: str, stream, sink) -> None:
    for line in stream:
        sink.write(f"[{prefix}] {line}")
        sink.flush()


def main(argv: Optional[Sequence[str]] = None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    cfg = parse_flags([a for a in argv if not a.startswith(("--job_name", "--task_id", "--task_index"))])
    passthrough = [a for a in argv if not a.startswith(("--job_name", "--task_id", "--task_index"))]
    procs: List = []
    for job, n in (("ps", cfg.n_ps), ("worker", cfg.n_workers)):
        for i in range(n):
            cmd = [sys.executable, "-m", "tensorflow_train_distributed_b200.train", *passthrough, f"--job_name={job}",
                   f"--task_id={i}"]
            p = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            t = threading.Thread(target=_pump, args=(f"{job}:{i}", p.stdout, sys.stdout), daemon=True)
            t.start()
            procs.append((job, i, p, t))
    rc = 0
    for job, i, p, t in procs:
        if job == "worker":
            rc = max(rc, abs(p.wait()))
    for job, i, p, t in procs:
        if job == "ps":
            try:
                p.wait(timeout=20)
            except subprocess.TimeoutExpired:
                p.terminate()
    for _, _, _, t in procs:
        t.join(timeout=2)
    return rc


if __name__ == "__main__":
    sys.exit(main())
