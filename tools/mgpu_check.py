"""Multi-GPU correctness check of the fused sync push (run under torchrun, one rank per GPU).

N replicas with batch B each, dropout off, must follow the same trajectory as ONE GPU training on the
concatenated batch of N*B examples (mean of the per-replica mean gradients == gradient of the global mean
loss) — i.e. K1 (peer reduce + update + broadcast of the copies) == reduce_scatter + update + all_gather.
Prints one JSON line from rank 0 and exits non-zero on mismatch.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


V = {}  # options of the variant being run (keys as the CHECK_* environment variables, which are the defaults)


def opt(key: str, default: str = "") -> str:
    return str(V.get(key, os.environ.get(key, default)))


def async_check(comm) -> int:
    """K2: Hogwild pushes. (a) atomic-sum exactness on integers-in-float, global_step counts every push;
    (b) a short training run converges and the shared global_step equals steps x workers."""
    from tensorflow_train_distributed_b200.config import TrainConfig
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    rank, world, B = comm.rank, comm.world_size, 128
    hosts = ",".join(f"127.0.0.1:{2222 + i}" for i in range(world))
    cfg = TrainConfig(worker_hosts=hosts, task_id=rank, batch_size=B, sync_replicas=False, save_checkpoint_secs=None,
                      seed=5, initial_learning_rate=1.0, learning_rate_decay_factor=1.0, regularizer="none",
                      use_moving_average=False, cuda_graph=False)
    tr = GpuTrainer(cfg, device_index=comm.device.index, comm=comm)
    tr.initialize()
    st = tr.store
    with torch.cuda.stream(tr.stream):
        st.w.zero_()
    tr.synchronize()
    comm.barrier()
    rounds = 3
    for _ in range(rounds):
        with torch.cuda.stream(tr.stream):
            st.grads.fill_(float(rank + 1))
            st.async_push()
        tr.synchronize()
    comm.barrier()
    expect = -float(rounds * sum(r + 1 for r in range(world)))
    own = st.w[st.lo:st.hi]
    valid = torch.zeros(st.layout.total, dtype=torch.bool, device=own.device)
    for g in st.layout.groups:
        valid[g.offset:g.offset + g.numel] = True
    exact = bool((own[valid[st.lo:st.hi]] == expect).all().item())
    gs_after_pushes = tr.read_values().global_step if rank == 0 else 0
    flag = torch.tensor([1 if exact else 0], device=comm.device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    tr.close()
    comm.barrier()

    steps = 60
    cfg2 = TrainConfig(worker_hosts=hosts, task_id=rank, batch_size=B, sync_replicas=False, save_checkpoint_secs=None, seed=9)
    tr2 = GpuTrainer(cfg2, device_index=comm.device.index, comm=comm)
    tr2.initialize()
    ds = synthetic_mnist(B * 64, seed=50 + rank)
    first = last = None
    for s in range(steps):
        xs, ys = ds.next_batch(B)
        last = tr2.train_step(xs, ys)
        first = first if first is not None else last.loss
    comm.barrier()
    final = tr2.read_values()
    gs0 = torch.tensor([final.global_step if rank == 0 else 0], device=comm.device)
    dist.all_reduce(gs0, op=dist.ReduceOp.MAX)
    vars_ = tr2.export_variables() if rank == 0 else None
    ok = True
    if rank == 0:
        finite = all(np.isfinite(v).all() for v in vars_.values())
        ok = bool(flag.item() == 1 and gs_after_pushes == rounds * world and int(gs0.item()) == steps * world
                  and last.loss < 0.7 * first and finite)
        print("MGPU " + json.dumps({"name": V.get("name", "async"), "mode": "async", "world": world, "atomic_sum_exact": bool(flag.item() == 1),
                                    "global_step_after_pushes": gs_after_pushes, "expected": rounds * world,
                                    "global_step_after_training": int(gs0.item()), "expected_training": steps * world,
                                    "first_loss": first, "last_loss": last.loss, "finite": finite, "ok": ok}), flush=True)
    okt = torch.tensor([1 if ok else 0], device=comm.device)
    dist.broadcast(okt, src=0)
    tr2.close()
    dist.barrier(device_ids=[comm.device.index])
    return 0 if int(okt.item()) == 1 else 1


def _common_cfg():
    """Model / optimiser variations shared by the sync and the restore check (all from the environment)."""
    model = opt("CHECK_MODEL", "mlp")
    common = dict(model=model, dropout_rate=0.0, save_checkpoint_secs=None, seed=77, sync_replicas=True,
                  activation="relu" if model == "wide_mlp" else "elu",
                  pull_fused=opt("CHECK_PULL_FUSED", "0") == "1",  # K3: W1 fetched inside the first GEMM
                  regularizer=opt("CHECK_REG", "l2"),
                  use_moving_average=opt("CHECK_EMA", "1") == "1",
                  step_impl=opt("CHECK_STEP_IMPL", "auto"))
    if opt("CHECK_HIDDEN"):
        common["hidden_units"] = [int(x) for x in opt("CHECK_HIDDEN").split(",")]
    return model, common


def restore_check(comm) -> int:
    """Save on the chief, tear every replica down, restore on EVERY replica (each GPU is the ps of one flat range,
    reference :208-215), continue: the continued trajectory must equal an uninterrupted run (dropout 0)."""
    import tempfile
    from tensorflow_train_distributed_b200.config import TrainConfig
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.io import bundle, checkpoint
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    rank, world = comm.rank, comm.world_size
    B = int(opt("CHECK_BATCH", "128"))
    s1, s2 = int(opt("CHECK_STEPS", "20")), int(opt("CHECK_STEPS2", "15"))
    model, common = _common_cfg()
    hosts = ",".join(f"127.0.0.1:{2222 + i}" for i in range(world))
    cfg = TrainConfig(worker_hosts=hosts, task_id=rank, batch_size=B, **common)
    F = cfg.input_dim

    def batches(n, skip=0):
        ds = synthetic_mnist(B * world * 8, input_dim=F, seed=123)
        for _ in range(skip):
            ds.next_batch(B * world)
        for _ in range(n):
            xs, ys = ds.next_batch(B * world)
            yield xs[rank * B:(rank + 1) * B], ys[rank * B:(rank + 1) * B]

    # uninterrupted run
    tr = GpuTrainer(cfg, device_index=comm.device.index, comm=comm)
    tr.initialize()
    for xs, ys in batches(s1 + s2):
        full = tr.train_step(xs, ys)
    want = tr.export_variables()
    tr.close()
    # interrupted run: s1 steps, checkpoint by the chief, new processes' worth of state, s2 more steps
    d = [opt("CHECK_DIR") or (tempfile.mkdtemp(prefix="ttd_restore_") if rank == 0 else None)]
    dist.broadcast_object_list(d, src=0)
    d = d[0]
    tr = GpuTrainer(cfg, device_index=comm.device.index, comm=comm)
    tr.initialize()
    for xs, ys in batches(s1):
        tr.train_step(xs, ys)
    snap = tr.export_variables() if rank == 0 else None  # chief-side peer reads: no collective
    if rank == 0:
        checkpoint.Saver(d, 5).save(snap, int(snap["global_step"]), {}, 1)
    tr.close()
    comm.barrier()
    restored = bundle.read_bundle(checkpoint.latest_checkpoint(d))  # EVERY rank reads the bundle
    tr = GpuTrainer(cfg.replace(seed=4242), device_index=comm.device.index, comm=comm)  # other seed: nothing may survive from init
    tr.initialize(restore=restored)
    gs0 = tr.read_values().global_step
    for xs, ys in batches(s2, skip=s1):
        cont = tr.train_step(xs, ys)
    got = tr.export_variables()
    gs_all = [torch.zeros(1, device=comm.device, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(gs_all, torch.tensor([tr.read_values().global_step], device=comm.device, dtype=torch.int64))
    ok = True
    if rank == 0:
        worst = 0.0
        for k, v in want.items():
            if k == "global_step":
                ok = ok and int(v) == int(got[k]) == s1 + s2
                continue
            worst = max(worst, float(np.abs(got[k] - v).max() / max(np.abs(v).max(), 1e-6)))
        same_gs = all(int(g.item()) == s1 + s2 for g in gs_all)
        ok = bool(ok and worst < 1e-5 and gs0 == s1 and same_gs and abs(cont.loss - full.loss) < 1e-5 * max(1.0, full.loss))
        print("MGPU " + json.dumps({"name": V.get("name", "restore"), "mode": "restore", "world": world, "model": model, "steps": [s1, s2],
                                    "global_step_after_restore": gs0, "global_steps_at_end": [int(g.item()) for g in gs_all],
                                    "worst_rel_param_err_vs_uninterrupted": worst, "loss_continued": cont.loss,
                                    "loss_uninterrupted": full.loss, "variables": len(got), "ok": ok}), flush=True)
    flag = torch.tensor([1 if ok else 0], device=comm.device)
    dist.broadcast(flag, src=0)
    tr.close()
    dist.barrier(device_ids=[comm.device.index])
    return 0 if int(flag.item()) == 1 else 1


def main() -> int:
    """One variant from the CHECK_* environment, or — CHECK_VARIANTS='[{"name": ..., "CHECK_MODE": ...}, ...]' — several
    in one process group (one NCCL / symmetric-memory bring-up instead of one per variant)."""
    from tensorflow_train_distributed_b200.parallel.symmetric import SymmetricComm
    comm = SymmetricComm.from_env()
    variants = json.loads(os.environ["CHECK_VARIANTS"]) if os.environ.get("CHECK_VARIANTS") else [{}]
    rc = 0
    for v in variants:
        V.clear()
        V.update({k: str(x) for k, x in v.items()})
        if "TTD_NVLS" in V:
            os.environ["TTD_NVLS"] = V["TTD_NVLS"]  # read by the parameter store when it is built
        else:
            os.environ.pop("TTD_NVLS", None)
        if comm.rank == 0 and V.get("name"):
            print(f"MGPU-VARIANT {V['name']}", flush=True)
        mode = opt("CHECK_MODE", "sync")
        one = async_check(comm) if mode == "async" else (restore_check(comm) if mode == "restore" else sync_check(comm))
        rc = rc or one
    dist.destroy_process_group()
    return rc


def sync_check(comm) -> int:
    from tensorflow_train_distributed_b200.config import TrainConfig
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    rank, world = comm.rank, comm.world_size
    steps, B = int(opt("CHECK_STEPS", "25")), int(opt("CHECK_BATCH", "128"))
    model, common = _common_cfg()
    hosts = ",".join(f"127.0.0.1:{2222 + i}" for i in range(world))
    cfg = TrainConfig(worker_hosts=hosts, task_id=rank, batch_size=B, **common)
    tr = GpuTrainer(cfg, device_index=comm.device.index, comm=comm)
    tr.initialize()
    F = cfg.input_dim
    ds = synthetic_mnist(B * world * 8, input_dim=F, seed=123)  # identical on every rank
    vals = None
    for s in range(steps):
        xs, ys = ds.next_batch(B * world)
        vals = tr.train_step(xs[rank * B:(rank + 1) * B], ys[rank * B:(rank + 1) * B])
    mine = tr.export_variables()  # gathers owner shards
    # replicas must hold bit-identical compute copies
    # (w_t[0] is owner-local under K3 — the first GEMM pulls it from the owners — so it is not compared)
    shared = tr.store.w_t[1:] + [t for t in tr.store.w_rm if t is not None]
    digest = torch.stack([t.float().sum() for t in shared] + [b.sum() for b in tr.store.bias_c]).double()
    all_d = [torch.zeros_like(digest) for _ in range(world)]
    dist.all_gather(all_d, digest)
    same_copies = all(torch.equal(all_d[0], d) for d in all_d)
    losses = [torch.zeros(1, device=comm.device, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(losses, torch.tensor([vals.loss], device=comm.device, dtype=torch.float64))
    mean_loss = float(torch.stack(losses).mean())
    ok, report = True, {}
    if rank == 0:
        ref = GpuTrainer(TrainConfig(worker_hosts="127.0.0.1:2222", batch_size=B * world, **common), device_index=comm.device.index)
        ref.initialize()
        ds2 = synthetic_mnist(B * world * 8, input_dim=F, seed=123)
        rv = None
        for s in range(steps):
            xs, ys = ds2.next_batch(B * world)
            rv = ref.train_step(xs, ys)
        want = ref.export_variables()
        worst = 0.0
        for k, v in want.items():
            if k == "global_step":
                ok = ok and int(v) == int(mine[k]) == steps
                continue
            denom = max(np.abs(v).max(), 1e-6)
            err = float(np.abs(mine[k] - v).max() / denom)
            worst = max(worst, err)
        ok = ok and worst < 2e-2 and same_copies and abs(mean_loss - rv.loss) < 5e-2 * max(1.0, rv.loss)
        report = {"name": V.get("name", "sync"), "mode": "sync", "world": world, "model": model, "steps": steps,
                  "nvls": bool(tr.store.use_multicast),
                  "variant": dict(V),
                  "worst_rel_param_err": worst, "same_copies": same_copies,
                  "mean_replica_loss": mean_loss, "single_gpu_loss": rv.loss, "global_step": int(mine["global_step"]), "ok": bool(ok)}
        print("MGPU " + json.dumps(report), flush=True)
    flag = torch.tensor([1 if ok else 0], device=comm.device)
    dist.broadcast(flag, src=0)
    tr.close()
    dist.barrier(device_ids=[comm.device.index])
    return 0 if int(flag.item()) == 1 else 1


if __name__ == "__main__":
    sys.exit(main())
