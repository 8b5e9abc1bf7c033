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
        print("MGPU " + json.dumps({"mode": "async", "world": world, "atomic_sum_exact": bool(flag.item() == 1),
                                    "global_step_after_pushes": gs_after_pushes, "expected": rounds * world,
                                    "global_step_after_training": int(gs0.item()), "expected_training": steps * world,
                                    "first_loss": first, "last_loss": last.loss, "finite": finite, "ok": ok}), flush=True)
    okt = torch.tensor([1 if ok else 0], device=comm.device)
    dist.broadcast(okt, src=0)
    tr2.close()
    dist.barrier(device_ids=[comm.device.index])
    dist.destroy_process_group()
    return 0 if int(okt.item()) == 1 else 1


def main() -> int:
    from tensorflow_train_distributed_b200.config import TrainConfig
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    from tensorflow_train_distributed_b200.parallel.symmetric import SymmetricComm
    comm = SymmetricComm.from_env()
    if os.environ.get("CHECK_MODE", "sync") == "async":
        return async_check(comm)
    rank, world = comm.rank, comm.world_size
    steps, B = int(os.environ.get("CHECK_STEPS", "25")), 128
    model = os.environ.get("CHECK_MODEL", "mlp")
    hosts = ",".join(f"127.0.0.1:{2222 + i}" for i in range(world))
    common = dict(model=model, dropout_rate=0.0, save_checkpoint_secs=None, seed=77, sync_replicas=True,
                  activation="relu" if model == "wide_mlp" else "elu",
                  pull_fused=os.environ.get("CHECK_PULL_FUSED", "0") == "1")  # K3: W1 fetched inside the first GEMM
    cfg = TrainConfig(worker_hosts=hosts, task_id=rank, batch_size=B, **common)
    tr = GpuTrainer(cfg, device_index=comm.device.index, comm=comm)
    tr.initialize()
    ds = synthetic_mnist(B * world * 8, seed=123)  # identical on every rank
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
        ds2 = synthetic_mnist(B * world * 8, seed=123)
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
        report = {"world": world, "model": model, "steps": steps, "worst_rel_param_err": worst, "same_copies": same_copies,
                  "mean_replica_loss": mean_loss, "single_gpu_loss": rv.loss, "global_step": int(mine["global_step"]), "ok": bool(ok)}
        print("MGPU " + json.dumps(report), flush=True)
    flag = torch.tensor([1 if ok else 0], device=comm.device)
    dist.broadcast(flag, src=0)
    tr.close()
    dist.barrier(device_ids=[comm.device.index])
    dist.destroy_process_group()
    return 0 if int(flag.item()) == 1 else 1


if __name__ == "__main__":
    sys.exit(main())
