"""Two more stock-PyTorch arms of the operative baseline. None of this
repository's kernels, stores or engine is on these paths.

``zero1``       the like-for-like comparison for the fused push/pull kernel (K1): gradients flattened into one
                buffer, NCCL ``reduce_scatter_tensor`` (mean), the owner rank applies SGD + L2 + EMA to its 1/N
                slice of an fp32 master, NCCL ``all_gather_into_tensor`` hands the refreshed parameters back.
``torch_graph`` single GPU: the whole eager step (bf16 autocast forward, backward, fused SGD, EMA lerp) captured in
                ONE CUDA graph and replayed — the strongest launch-latency-free form of the stock stack. The
                learning rate is a device tensor so that the schedule survives capture.

Model / data / timing are those of ``baseline/ddp_mlp.py``.
"""
from __future__ import annotations

import json
import os

import torch
import torch.distributed as dist
import torch.nn as nn

from .ddp_mlp import build_model

SIZES = {"mnist_sync": [200, 100, 50, 25, 10], "mnist_async": [200, 100, 50, 25, 10], "wide_sync": [8192, 8192, 10]}


def _timed(fn, steps, world, local_rank, device):
    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def _data(args, rank, device):
    B, F = args.batch, 784
    pool = max(2, -(-160 * 2**20 // (B * F * 4)))
    g = torch.Generator(device=device).manual_seed(100 + rank)
    xs = torch.rand(pool, B, F, device=device, generator=g)
    ys = torch.randint(0, 10, (pool, B), device=device, generator=g)
    return B, F, pool, xs, ys


def _emit(impl, args, world, ms, ms_e2e, sizes, B, F, extra):
    ex = B * world
    out = {"metric": "examples/sec", "value": ex * args.steps / (ms / 1e3), "unit": "examples/s", "n_gpus": world,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "bf16 autocast", "impl": impl,
           "data": "synthetic MNIST-shaped, random-init weights",
           "config": dict({"model": "-".join(map(str, [F] + sizes)), "global_batch": ex}, **extra)}
    if ms_e2e is not None:
        out["e2e"] = {"value": ex * args.steps / (ms_e2e / 1e3), "unit": "examples/s", "ms_per_step": ms_e2e / args.steps,
                      "h2d_bytes_per_step": B * F * 4 + B * 8, "d2h_bytes_per_step": 4}
    print(json.dumps(out))


def run_zero1_bench(args, rank, world, local_rank) -> int:
    device = torch.device("cuda", local_rank)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    sizes = SIZES.get(args.config)
    if sizes is None:
        if rank == 0:
            print(json.dumps({"impl": "zero1", "unavailable": f"no baseline for config {args.config}"}))
        return 0
    B, F, pool, xs, ys = _data(args, rank, device)
    torch.manual_seed(1234)
    model = build_model(F, sizes).to(device)
    params = list(model.parameters())
    total = sum(p.numel() for p in params)
    shard = -(-total // world)
    padded = shard * world
    # one flat fp32 buffer for parameters and one for gradients; the module's tensors become views of them
    flat_p = torch.zeros(padded, device=device)
    flat_g = torch.zeros(padded, device=device)
    reg = torch.zeros(padded, device=device)  # 1 where L2 applies (hidden kernels only)
    lin = [m for m in model if isinstance(m, nn.Linear)]
    off = 0
    for p in params:
        n = p.numel()
        flat_p[off:off + n].copy_(p.detach().reshape(-1))
        p.data = flat_p[off:off + n].view_as(p)
        p.grad = flat_g[off:off + n].view_as(p)
        if any(p is m.weight for m in lin[:-1]):
            reg[off:off + n] = 1.0
        off += n
    lo = rank * shard
    my_p = flat_p[lo:lo + shard]
    my_reg = reg[lo:lo + shard]
    my_ema = my_p.clone()
    my_g = torch.zeros(shard, device=device)
    step_no = [0]

    def step(x, y):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = model(x)
        loss = nn.functional.cross_entropy(logits.float(), y)
        flat_g.zero_()
        loss.backward()  # accumulates into the views of flat_g
        if world > 1:
            dist.reduce_scatter_tensor(my_g, flat_g, op=dist.ReduceOp.AVG)
            g = my_g
        else:
            g = flat_g[:shard]
        n = step_no[0]
        lr = 0.01 * 0.96 ** (n // 468)
        my_p.add_(g + 0.01 * my_reg * my_p, alpha=-lr)
        d = min(0.99, (2 + n) / (11 + n))
        my_ema.lerp_(my_p, 1 - d)
        if world > 1:
            dist.all_gather_into_tensor(flat_p, my_p)
        step_no[0] = n + 1
        return loss

    def dev_step(i):
        step(xs[i % pool], ys[i % pool])

    for i in range(max(args.warmup, 3)):
        dev_step(i)
    ms = _timed(dev_step, args.steps, world, local_rank, device)
    if rank == 0:
        _emit("zero1", args, world, ms, None, sizes, B, F,
              {"parallelism": f"dp{world} NCCL reduce_scatter -> sharded SGD+L2+EMA -> all_gather (eager)"})
    if world > 1:
        dist.barrier(device_ids=[local_rank])
        dist.destroy_process_group()
    return 0


def run_torch_graph_bench(args, rank, world, local_rank) -> int:
    if world != 1:
        if rank == 0:
            print(json.dumps({"impl": "torch_graph", "unavailable": "single-GPU arm (whole-step CUDA graph of stock PyTorch)"}))
        return 0
    device = torch.device("cuda", local_rank)
    sizes = SIZES.get(args.config)
    if sizes is None:
        print(json.dumps({"impl": "torch_graph", "unavailable": f"no baseline for config {args.config}"}))
        return 0
    B, F, pool, xs, ys = _data(args, rank, device)
    torch.manual_seed(1234)
    model = build_model(F, sizes).to(device)
    params = list(model.parameters())
    lin = [m for m in model if isinstance(m, nn.Linear)]
    hidden_w = [m.weight for m in lin[:-1]]
    ema = [p.detach().clone() for p in params]
    lr_t = torch.full((), 0.01, device=device)          # device-side schedule state: survives graph capture
    n_t = torch.zeros((), device=device)
    x_static = torch.zeros(B, F, device=device)
    y_static = torch.zeros(B, device=device, dtype=torch.int64)
    loss_out = torch.zeros((), device=device)

    def step_body():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = model(x_static)
        loss = nn.functional.cross_entropy(logits.float(), y_static)
        grads = torch.autograd.grad(loss, params)
        with torch.no_grad():
            reg = [0.01 * w for w in hidden_w]
            torch._foreach_add_([g for g, p in zip(grads, params) if any(p is w for w in hidden_w)], reg)
            upd = torch._foreach_mul(list(grads), lr_t)
            torch._foreach_sub_(params, upd)
            n_t.add_(1)
            d = torch.minimum(torch.full_like(n_t, 0.99), (1 + n_t) / (10 + n_t))
            w = 1 - d
            for e, p in zip(ema, params):  # tensor weight: no host read-back inside the capture
                e.lerp_(p, w)
            lr_t.copy_(0.01 * torch.pow(torch.full_like(n_t, 0.96), torch.floor(n_t / 468)))
            loss_out.copy_(loss.detach())

    side = torch.cuda.Stream(device=device)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            step_body()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step_body()

    def dev_step(i):
        j = i % pool
        x_static.copy_(xs[j], non_blocking=True)   # device -> device: keeps the input larger than L2 over time
        y_static.copy_(ys[j], non_blocking=True)
        graph.replay()

    hx, hy = xs[:64].cpu().pin_memory(), ys[:64].cpu().pin_memory()
    loss_host = torch.zeros(64, dtype=torch.float32).pin_memory()

    def e2e_step(i):
        j = i % 64
        x_static.copy_(hx[j], non_blocking=True)
        y_static.copy_(hy[j], non_blocking=True)
        graph.replay()
        loss_host[j:j + 1].copy_(loss_out.reshape(1), non_blocking=True)

    for i in range(max(args.warmup, 3)):
        dev_step(i)
    ms = _timed(dev_step, args.steps, 1, local_rank, device)
    for i in range(max(args.warmup, 3)):
        e2e_step(i)
    ms_e2e = _timed(e2e_step, args.steps, 1, local_rank, device)
    _emit("torch_graph", args, 1, ms, ms_e2e, sizes, B, F,
          {"parallelism": "dp1, whole step (autocast fwd, bwd, SGD+L2, EMA, lr schedule) as one CUDA graph of stock PyTorch ops",
           "final_loss": float(loss_out)})
    return 0
