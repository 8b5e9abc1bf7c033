"""Operative baseline: the identical model/config in stock PyTorch — DistributedDataParallel
over NCCL, cuBLAS GEMMs, bf16 autocast, fused SGD — on the same synthetic data. None of this repository's
kernels, stores or engine is on this path.

Model: distribute_training.py:39-63 (784-200-100-50-25-10, ELU, dropout 0.01), loss :77-82, accuracy :96/:125,
lr schedule :136-140, SGD :150, L2 on hidden kernels and EMA per the north-star.
"""
from __future__ import annotations

import json
import os

import torch
import torch.distributed as dist
import torch.nn as nn


def build_model(input_dim, sizes, act="elu", drop=0.01):
    layers, fan_in = [], input_dim
    for i, n in enumerate(sizes):
        lin = nn.Linear(fan_in, n)
        nn.init.trunc_normal_(lin.weight, std=(2.6 / fan_in) ** 0.5, a=-2 * (2.6 / fan_in) ** 0.5, b=2 * (2.6 / fan_in) ** 0.5)
        nn.init.zeros_(lin.bias)
        layers.append(lin)
        if i < len(sizes) - 1:
            layers.append(nn.ELU() if act == "elu" else nn.ReLU())
            layers.append(nn.Dropout(drop))
        fan_in = n
    return nn.Sequential(*layers)


def run_ddp_bench(args, rank, world, local_rank, keep_group: bool = False) -> int:
    device = torch.device("cuda", local_rank)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    sizes = {"mnist_sync": [200, 100, 50, 25, 10], "mnist_async": [200, 100, 50, 25, 10],
             "wide_sync": [8192, 8192, 10]}.get(args.config)
    if sizes is None:
        if rank == 0:
            print(json.dumps({"impl": "ddp", "unavailable": f"no DDP baseline for config {args.config}"}))
        return 0
    B, F = args.batch, 784
    torch.manual_seed(1234)
    model = build_model(F, sizes).to(device)
    hidden_w = [m.weight for m in model if isinstance(m, nn.Linear)][:-1]
    other = [p for p in model.parameters() if all(p is not q for q in hidden_w)]
    opt = torch.optim.SGD([{"params": hidden_w, "weight_decay": 0.01}, {"params": other, "weight_decay": 0.0}],
                          lr=0.01, foreach=True)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 0.96 ** (s // 468))
    ema = [p.detach().clone() for p in model.parameters()]
    net = nn.parallel.DistributedDataParallel(model, device_ids=[local_rank]) if world > 1 else model
    pool = max(2, -(-160 * 2**20 // (B * F * 4)))
    g = torch.Generator(device=device).manual_seed(100 + rank)
    xs = torch.rand(pool, B, F, device=device, generator=g)
    ys = torch.randint(0, 10, (pool, B), device=device, generator=g)
    hx, hy = xs[:64].cpu().pin_memory(), ys[:64].cpu().pin_memory()
    x_static = torch.zeros(B, F, device=device)
    y_static = torch.zeros(B, device=device, dtype=torch.int64)
    loss_host = torch.zeros(64, dtype=torch.float32).pin_memory()
    step_no = [0]

    def step(x, y):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = net(x)
        loss = nn.functional.cross_entropy(logits.float(), y)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        sched.step()
        n = step_no[0] + 1
        d = min(0.99, (1 + n) / (10 + n))
        torch._foreach_lerp_(ema, [p.detach() for p in model.parameters()], 1 - d)
        step_no[0] = n
        return loss

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize()

    def timed(fn, steps):
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

    def dev_step(i):
        step(xs[i % pool], ys[i % pool])

    def e2e_step(i):
        j = i % 64
        x_static.copy_(hx[j], non_blocking=True)
        y_static.copy_(hy[j], non_blocking=True)
        loss = step(x_static, y_static)
        loss_host[j:j + 1].copy_(loss.detach().reshape(1), non_blocking=True)

    for i in range(max(args.warmup, 3)):
        dev_step(i)
    ms = timed(dev_step, args.steps)
    for i in range(max(args.warmup, 3)):
        e2e_step(i)
    ms_e2e = timed(e2e_step, args.steps)
    if rank == 0:
        ex = B * world
        print(json.dumps({
            "metric": "examples/sec", "value": ex * args.steps / (ms / 1e3), "unit": "examples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16 autocast", "impl": "ddp",
            "data": "synthetic MNIST-shaped, random-init weights",
            "config": {"model": "-".join(map(str, [F] + sizes)), "global_batch": ex, "parallelism": f"dp{world} DDP/NCCL"},
            "e2e": {"value": ex * args.steps / (ms_e2e / 1e3), "unit": "examples/s", "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": B * F * 4 + B * 8, "d2h_bytes_per_step": 4},
        }))
    if world > 1 and not keep_group:
        dist.barrier(device_ids=[local_rank])
        dist.destroy_process_group()
    return 0
