"""BASELINE config 1: real ps + worker processes on localhost over gloo (CPU tier, multi-process)."""
import os
import re
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def _launch(n_ps, n_workers, extra, ckpt_dir, timeout=240):
    ports = _free_ports(n_ps + n_workers)
    ps_hosts = ",".join(f"127.0.0.1:{p}" for p in ports[:n_ps])
    w_hosts = ",".join(f"127.0.0.1:{p}" for p in ports[n_ps:])
    base = [sys.executable, "-m", "tensorflow_train_distributed_b200.train", f"--ps_hosts={ps_hosts}",
            f"--worker_hosts={w_hosts}", "--device=cpu", "--dataset_size=4096", f"--checkpoint_dir={ckpt_dir}",
            "--log_every_n_steps=50"] + extra
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="1", MKL_NUM_THREADS="1")
    procs = []
    for i in range(n_ps):
        procs.append(("ps", i, subprocess.Popen(base + ["--job_name=ps", f"--task_id={i}"], env=env, cwd=ROOT,
                                                stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for i in range(n_workers):
        procs.append(("worker", i, subprocess.Popen(base + ["--job_name=worker", f"--task_id={i}"], env=env, cwd=ROOT,
                                                    stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    outs = {}
    try:
        for job, i, p in sorted(procs, key=lambda t: t[0] != "worker"):  # workers first: a ps outlives them
            out, _ = p.communicate(timeout=timeout if job == "worker" else 30)
            outs[(job, i)] = (p.returncode, out)
    finally:
        for _, _, p in procs:
            if p.poll() is None:
                p.kill()
    return outs


def _final(out):
    m = re.search(r"total step: (\d+), global_step: (\d+)", out)
    assert m, out[-2000:]
    return int(m.group(1)), int(m.group(2))


def _losses(out):
    return [float(x) for x in re.findall(r"loss on training batch is ([0-9.e+-]+)", out)]


@pytest.mark.timeout(300)
@pytest.mark.parametrize("transport", ["gloo", "tcp"])
def test_async_one_ps_two_workers(tmp_path, transport):
    """The reference's default topology (:26, :29): the 300 global steps are SHARED by the workers (:220-222)."""
    d = str(tmp_path / "ck")
    outs = _launch(1, 2, ["--training_steps=300", "--save_checkpoint_secs=2", "--prefetch=2", f"--transport={transport}"], d)
    for k, (rc, out) in outs.items():
        assert rc == 0, (k, out[-3000:])
    (s0, g0), (s1, g1) = _final(outs[("worker", 0)][1]), _final(outs[("worker", 1)][1])
    assert g0 >= 300 and g1 >= 300
    assert 300 <= s0 + s1 <= 304  # the global steps are SHARED: together the workers run ~300 local steps, not 600
    assert "session started" in outs[("worker", 0)][1]
    l0 = _losses(outs[("worker", 0)][1])
    assert l0 and l0[-1] < 1.0  # learned something from 2.30
    from tensorflow_train_distributed_b200.io import bundle, checkpoint
    prefix = checkpoint.latest_checkpoint(d)
    assert prefix is not None
    t = bundle.read_bundle(prefix)
    assert int(t["global_step"]) >= 300 and t["hidden1/kernel"].shape == (784, 200)
    assert len(t) == 21  # 11 reference entries + 10 EMA shadows
    assert "applied" in outs[("ps", 0)][1]


@pytest.mark.timeout(300)
@pytest.mark.parametrize("transport", ["gloo", "tcp"])
def test_sync_replicas_two_ps_two_workers(tmp_path, transport):
    """SyncReplicasOptimizer semantics (A.2): every worker runs ~last_step local steps, one global step per
    aggregate, stale second gradient dropped, variables spread round-robin over two ps tasks."""
    d = str(tmp_path / "ck")
    outs = _launch(2, 2, ["--training_steps=120", "--sync_replicas", "--save_checkpoint_secs=none",
                          "--save_checkpoint_steps=60", "--dropout_rate=0", f"--transport={transport}"], d)
    for k, (rc, out) in outs.items():
        assert rc == 0, (k, out[-3000:])
    (s0, g0), (s1, g1) = _final(outs[("worker", 0)][1]), _final(outs[("worker", 1)][1])
    assert g0 >= 120 and g1 >= 120
    # every global step consumes two gradients; tokens are not bound to a worker, so one replica may run a few
    # local steps more than the other (same in SyncReplicasOptimizer), but together they cover 2 x 120
    assert s0 + s1 >= 240
    if transport == "gloo":  # the process-group rendezvous starts both workers together: near lock-step
        assert abs(s0 - 120) <= 8 and abs(s1 - 120) <= 8
    # tcp: as with TF, nothing makes a worker wait for its peers — tokens are not bound to a worker, so one that
    # starts earlier can contribute both gradients of an aggregate until the other has connected
    ps0 = outs[("ps", 0)][1]
    m = re.search(r"'applied': (\d+), 'dropped_stale': (\d+)", ps0)
    assert m and int(m.group(1)) >= 120 and int(m.group(2)) >= 1
    from tensorflow_train_distributed_b200.io import bundle, checkpoint
    r = bundle.BundleReader(checkpoint.latest_checkpoint(d))
    assert r.num_shards == 2 and r.entries["global_step"]["shard_id"] == 0 and r.entries["hidden1/kernel"]["shard_id"] == 1


@pytest.mark.timeout(400)
@pytest.mark.parametrize("transport", ["gloo", "tcp"])
def test_kill_and_resume_from_checkpoint(tmp_path, transport):
    """--inject_fault kills the chief mid-run; a relaunch on the same --checkpoint_dir resumes (:208)."""
    d = str(tmp_path / "ck")
    common = ["--training_steps=200", "--save_checkpoint_secs=none", "--save_checkpoint_steps=40", f"--transport={transport}"]
    outs = _launch(1, 1, common + ["--inject_fault=step=90,rank=0", "--ps_idle_timeout_secs=3"], d, timeout=120)
    assert "idle_timeout" in outs[("ps", 0)][1]  # the ps noticed that its only worker is gone
    assert outs[("worker", 0)][0] != 0  # killed
    from tensorflow_train_distributed_b200.io import checkpoint
    mid = checkpoint.checkpoint_step(checkpoint.latest_checkpoint(d))
    assert mid is not None and 40 <= mid < 200
    outs = _launch(1, 1, common, d)
    rc, out = outs[("worker", 0)]
    assert rc == 0, out[-3000:]
    steps, gs = _final(out)
    assert gs >= 200 and steps <= 200 - mid + 2  # only the remaining steps were run


@pytest.mark.timeout(300)
def test_native_ps_survives_a_worker_restart(tmp_path):
    """tcp transport: the ps is a listening server that keeps its state when its worker dies; the restarted
    chief reconnects, restores the last checkpoint into the live ps and finishes (the recovery the reference
    gets from MonitoredTrainingSession + a long-lived tf.train.Server, :175-181, :208)."""
    d = str(tmp_path / "ck")
    ports = _free_ports(2)
    base = [sys.executable, "-m", "tensorflow_train_distributed_b200.train", f"--ps_hosts=127.0.0.1:{ports[0]}",
            f"--worker_hosts=127.0.0.1:{ports[1]}", "--device=cpu", "--dataset_size=4096", f"--checkpoint_dir={d}",
            "--log_every_n_steps=50", "--transport=tcp", "--training_steps=200", "--save_checkpoint_secs=none",
            "--save_checkpoint_steps=40"]
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="1", MKL_NUM_THREADS="1")
    ps = subprocess.Popen(base + ["--job_name=ps", "--task_id=0"], env=env, cwd=ROOT, stdout=subprocess.PIPE,
                          stderr=subprocess.STDOUT, text=True)
    try:
        w1 = subprocess.run(base + ["--job_name=worker", "--task_id=0", "--inject_fault=step=90,rank=0"], env=env,
                            cwd=ROOT, capture_output=True, text=True, timeout=120)
        assert w1.returncode != 0 and ps.poll() is None  # the worker was killed, the ps is still serving
        w2 = subprocess.run(base + ["--job_name=worker", "--task_id=0"], env=env, cwd=ROOT, capture_output=True,
                            text=True, timeout=120)
        assert w2.returncode == 0, (w2.stdout + w2.stderr)[-3000:]
        steps, gs = _final(w2.stdout + w2.stderr)
        assert gs >= 200 and steps <= 200 - 40 + 2  # resumed from a checkpoint at step >= 40
        out, _ = ps.communicate(timeout=30)
        assert ps.returncode == 0 and "ps 0 done" in out
    finally:
        if ps.poll() is None:
            ps.kill()


def test_cpu_worker_model_covers_lenet():
    """Config 5 on the CPU tier: the worker-side model handles the LeNet layout (HWIO conv kernels, SAME conv1,
    TF flatten order) and its gradients have the variables' shapes."""
    import torch
    from tensorflow_train_distributed_b200.parallel import sharding
    from tensorflow_train_distributed_b200.parallel.cpu_ps import CpuMlp, he_init
    lay = sharding.ParamLayout(sharding.lenet_groups(10))
    p = {k: torch.from_numpy(v) for k, v in he_init(lay, 1).items()}
    assert sum(v.numel() for v in p.values()) == 61706
    m = CpuMlp(lay, "elu", 0.0, 3)
    g = torch.Generator().manual_seed(0)
    x, y = torch.rand(32, 784, generator=g), torch.randint(0, 10, (32,), generator=g)
    first = None
    for _ in range(40):
        loss, acc, grads = m.loss_and_grads(p, x, y)
        first = first or loss
        assert all(grads[k].shape == p[k].shape for k in p)
        p = {k: p[k] - 0.05 * grads[k] for k in p}
    assert loss < 0.3 * first and acc > 0.9


@pytest.mark.timeout(300)
def test_cluster_launcher_spawns_every_task(tmp_path):
    """``python -m ...launch``: one subprocess per ps / worker task of the ClusterSpec, prefixed output, exit code."""
    ports = _free_ports(3)
    cmd = [sys.executable, "-m", "tensorflow_train_distributed_b200.launch", f"--ps_hosts=127.0.0.1:{ports[0]}",
           f"--worker_hosts=127.0.0.1:{ports[1]},127.0.0.1:{ports[2]}", "--device=cpu", "--transport=tcp",
           "--dataset_size=2048", "--training_steps=80", "--save_checkpoint_secs=none", "--log_every_n_steps=20",
           f"--checkpoint_dir={tmp_path / 'ck'}"]
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="1", MKL_NUM_THREADS="1")
    r = subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    out = r.stdout
    assert "[ps:0] " in out and "[worker:0] " in out and "[worker:1] " in out
    finals = re.findall(r"\[worker:(\d)\] .*total step: (\d+), global_step: (\d+)", out)
    assert len(finals) == 2 and all(int(g) >= 80 for _, _, g in finals)
    assert "ps 0 done" in out


@pytest.mark.timeout(300)
def test_launcher_restarts_the_group_after_a_worker_dies(tmp_path):
    """``--max_restarts``: the launcher is the supervisor of the recoverable session (:208). Worker 1 is killed at
    its 60th local step; the group is stopped, started again, restores the newest checkpoint and finishes."""
    ports = _free_ports(3)
    d = tmp_path / "ck"
    cmd = [sys.executable, "-m", "tensorflow_train_distributed_b200.launch", f"--ps_hosts=127.0.0.1:{ports[0]}",
           f"--worker_hosts=127.0.0.1:{ports[1]},127.0.0.1:{ports[2]}", "--device=cpu", "--transport=tcp", "--sync_replicas",
           "--dataset_size=2048", "--training_steps=150", "--save_checkpoint_secs=none", "--save_checkpoint_steps=25",
           "--log_every_n_steps=20", f"--checkpoint_dir={d}", "--max_restarts=1", "--inject_fault=step=60,rank=1",
           "--ps_idle_timeout_secs=20"]
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="1", MKL_NUM_THREADS="1")
    r = subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True, timeout=280)
    out = r.stdout
    assert r.returncode == 0, (out + r.stderr)[-3000:]
    assert "restarting the group from the newest checkpoint (restart 1 of 1)" in out
    finals = re.findall(r"\[worker:(\d)\] .*total step: (\d+), global_step: (\d+)", out)
    assert len(finals) == 2 and all(int(g) >= 150 for _, _, g in finals), finals
    # the second run resumed (its local step count is what was left, not the whole 150)
    assert all(int(s) < 150 for _, s, _ in finals), finals


@pytest.mark.timeout(300)
def test_evaluate_cli_scores_a_trained_checkpoint(tmp_path):
    """Train on the CPU tier, then ``python -m ...evaluate``: dropout off, weights or moving averages, same data set."""
    d = str(tmp_path / "ck")
    outs = _launch(1, 2, ["--training_steps=300", "--save_checkpoint_secs=none", "--save_checkpoint_steps=100", "--transport=tcp"], d)
    for k, (rc, out) in outs.items():
        assert rc == 0, (k, out[-3000:])
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="1")
    for extra in ([], ["--ema"]):
        r = subprocess.run([sys.executable, "-m", "tensorflow_train_distributed_b200.evaluate", f"--checkpoint_dir={d}",
                            "--dataset_size=4096"] + extra, env=env, cwd=ROOT, capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, (r.stdout + r.stderr)[-2000:]
        m = re.search(r"global_step (\d+), (\d+) examples, loss ([0-9.e+-]+), accuracy ([0-9.]+)", r.stdout)
        assert m, r.stdout
        assert int(m.group(1)) >= 300 and int(m.group(2)) == 4096
        assert float(m.group(3)) < 0.5 and float(m.group(4)) > 0.95
