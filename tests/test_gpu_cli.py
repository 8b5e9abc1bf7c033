"""The user-facing path on GPUs: `launch` -> `train` -> MonitoredTrainingSession -> GpuTrainer (GPU tier)."""
import os
import re
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    out = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return out


def _run(n_workers, extra, ckpt, timeout=300):
    ports = _ports(n_workers + 1)
    cmd = [sys.executable, "-m", "tensorflow_train_distributed_b200.launch", f"--ps_hosts=127.0.0.1:{ports[0]}",
           "--worker_hosts=" + ",".join(f"127.0.0.1:{p}" for p in ports[1:]), f"--checkpoint_dir={ckpt}",
           "--dataset_size=8192", "--log_every_n_steps=100"] + extra
    p = subprocess.run(cmd, cwd=ROOT, env=dict(os.environ, PYTHONPATH=ROOT), capture_output=True, text=True, timeout=timeout)
    return p.returncode, p.stdout + p.stderr


def _finals(out):
    return [(int(a), int(b)) for a, b in re.findall(r"total step: (\d+), global_step: (\d+)", out)]


def test_single_gpu_cli_trains_checkpoints_and_resumes(tmp_path):
    d = str(tmp_path / "ck")
    rc, out = _run(1, ["--training_steps=400", "--save_checkpoint_secs=none", "--save_checkpoint_steps=150"], d)
    assert rc == 0, out[-3000:]
    assert "session started" in out and _finals(out) == [(400, 400)]
    losses = [float(x) for x in re.findall(r"loss on training batch is ([0-9.e+-]+)", out)]
    assert losses and losses[-1] < 0.5
    from tensorflow_train_distributed_b200.io import bundle, checkpoint
    assert checkpoint.read_state(d)["all_model_checkpoint_paths"] == ["model.ckpt-0", "model.ckpt-150", "model.ckpt-300", "model.ckpt-400"]
    t = bundle.read_bundle(checkpoint.latest_checkpoint(d))
    assert int(t["global_step"]) == 400 and len(t) == 21
    rc, out = _run(1, ["--training_steps=450", "--save_checkpoint_secs=none", "--save_checkpoint_steps=1000"], d)
    assert rc == 0, out[-3000:]
    assert _finals(out) == [(50, 450)]  # resumed at 400 (:208), ran only the remaining 50 steps


@pytest.mark.multigpu
def test_two_gpu_sync_replicas_cli(tmp_path):
    d = str(tmp_path / "ck")
    rc, out = _run(2, ["--training_steps=300", "--sync_replicas", "--save_checkpoint_secs=1"], d)
    assert rc == 0, out[-3000:]
    assert _finals(out) == [(300, 300), (300, 300)]  # sync: every worker runs all the steps
    from tensorflow_train_distributed_b200.io import bundle, checkpoint
    t = bundle.read_bundle(checkpoint.latest_checkpoint(d))
    assert int(t["global_step"]) == 300 and t["hidden1/kernel"].shape == (784, 200)


@pytest.mark.multigpu
def test_two_gpu_async_replicas_cli(tmp_path):
    d = str(tmp_path / "ck")
    rc, out = _run(2, ["--training_steps=400", "--save_checkpoint_secs=none", "--save_checkpoint_steps=200"], d)
    assert rc == 0, out[-3000:]
    fin = _finals(out)
    assert len(fin) == 2 and all(g >= 400 for _, g in fin)
    assert 400 <= sum(s for s, _ in fin) <= 408  # the global steps are shared (:220-222): ~400 local steps in total


@pytest.mark.multigpu
def test_two_gpu_sync_pull_fused_cli(tmp_path):
    """K3 (``--pull_fused``): the first GEMM fetches W1 from its owners by TMA instead of K1 broadcasting it."""
    d = str(tmp_path / "ck")
    rc, out = _run(2, ["--training_steps=300", "--sync_replicas", "--pull_fused", "--save_checkpoint_secs=none"], d)
    assert rc == 0, out[-3000:]
    assert _finals(out) == [(300, 300), (300, 300)]
    losses = [float(x) for x in re.findall(r"loss on training batch is ([0-9.e+-]+)", out)]
    assert losses and losses[-1] < 0.5
