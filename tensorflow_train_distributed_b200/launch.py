"""Spawn a whole cluster on this machine — what the reference's operator does by hand with one terminal per
``python distribute_training.py --job_name=... --task_id=...`` (README.md, flag defaults at :24-36).

    python -m tensorflow_train_distributed_b200.launch --ps_hosts=127.0.0.1:2221 \
        --worker_hosts=127.0.0.1:2222,127.0.0.1:2223 [any training flag ...]

Every ps and worker task of the cluster becomes one subprocess of ``...train`` with its ``--job_name`` /
``--task_id``; worker output is prefixed with its task. The exit code is the worst worker exit code; ps tasks
that outlive the workers (the reference's never exit, :181) are terminated once the workers are done.

``--max_restarts=N`` makes the launcher the supervisor of a recoverable session (the reference relies on
``_RecoverableSession`` inside ``MonitoredTrainingSession``, :208): when a worker dies, the surviving tasks are
stopped (on GPUs they would otherwise wait for the dead replica in the next update's barrier until its watchdog
fires), and the whole group is started again; every worker then restores the newest checkpoint of the — pinned —
checkpoint directory and training carries on from that global step.
"""
from __future__ import annotations

import subprocess
import sys
import threading
import time
from typing import List, Optional, Sequence

from .config import parse_flags

_ROLE_FLAGS = ("--job_name", "--task_id", "--task_index")


def _pump(prefix: str, stream, sink) -> None:
    for line in stream:
        sink.write(f"[{prefix}] {line}")
        sink.flush()


def _spawn(cfg, passthrough: List[str]) -> List:
    procs: List = []
    for job, n in (("ps", cfg.n_ps), ("worker", cfg.n_workers)):
        for i in range(n):
            cmd = [sys.executable, "-m", "tensorflow_train_distributed_b200.train", *passthrough, f"--job_name={job}",
                   f"--task_id={i}"]
            p = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            t = threading.Thread(target=_pump, args=(f"{job}:{i}", p.stdout, sys.stdout), daemon=True)
            t.start()
            procs.append((job, i, p, t))
    return procs


def _wait_workers(procs: List, stop_on_failure: bool) -> int:
    """Worst worker exit code. With ``stop_on_failure`` the first failing worker ends the wait early."""
    workers = [p for job, _, p, _ in procs if job == "worker"]
    rc = 0
    while True:
        alive = False
        for p in workers:
            r = p.poll()
            if r is None:
                alive = True
            elif r != 0:
                rc = max(rc, abs(r))
        if not alive or (rc != 0 and stop_on_failure):
            return rc
        time.sleep(0.05)


def _stop_all(procs: List, grace: float = 20.0) -> None:
    for job, _, p, _ in procs:
        if p.poll() is None and job == "ps":
            try:
                p.wait(timeout=grace)
            except subprocess.TimeoutExpired:
                p.terminate()
    for _, _, p, _ in procs:
        if p.poll() is None:
            p.terminate()
    for _, _, p, t in procs:
        try:
            p.wait(timeout=10)
        except subprocess.TimeoutExpired:
            p.kill()
        t.join(timeout=2)


def main(argv: Optional[Sequence[str]] = None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    passthrough = [a for a in argv if not a.startswith(_ROLE_FLAGS)]
    cfg = parse_flags(passthrough)
    if cfg.max_restarts > 0 and not cfg.checkpoint_dir:
        # a restarted group must find the directory of the run it continues: pin the time-stamped default now
        passthrough.append(f"--checkpoint_dir={cfg.resolve_checkpoint_dir()}")
    attempt = 0
    while True:
        # an injected fault is a one-shot test hook: the restarted group must not die at the same step again
        args = [a for a in passthrough if attempt == 0 or not a.startswith("--inject_fault")]
        procs = _spawn(cfg, args)
        rc = _wait_workers(procs, stop_on_failure=cfg.max_restarts > 0)
        if rc == 0 or attempt >= cfg.max_restarts:
            _stop_all(procs)
            return rc
        attempt += 1
        sys.stdout.write(f"[launch] a worker exited with code {rc}; restarting the group from the newest checkpoint "
                         f"(restart {attempt} of {cfg.max_restarts})\n")
        sys.stdout.flush()
        _stop_all(procs, grace=0.0)


if __name__ == "__main__":
    sys.exit(main())
