#!/usr/bin/env python
"""Headline benchmark: examples/sec of the reference training step on N B200 GPUs of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference|ddp|zero1|torch_graph] [--config mnist_sync|...]

Metric and config are BASELINE.json's: the distribute_training.py model (784-200-100-50-25-10 ELU MLP, dropout
0.01, batch 128 per worker, SGD with staircase-decayed lr), synchronous replicas on a GPU-sharded parameter
server, synthetic MNIST-shaped data, random-init weights. Weak scaling: every GPU processes 128 examples per
step; `value` is the whole job's examples/sec.

For N > 1 launch with torchrun (one rank per GPU); rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (model, sync, extra cfg)
    "mnist_sync": dict(model="mlp", sync_replicas=True),
    "mnist_async": dict(model="mlp", sync_replicas=False),
    "wide_sync": dict(model="wide_mlp", sync_replicas=True),
    "lenet_sync": dict(model="lenet", sync_replicas=True),
}


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (NVML, 5 ms period)."""

    REASONS = {
        0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x10: "sync_boost",
        0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
        0x100: "display_clock_setting",
    }

    def __init__(self, index: int):
        self.index, self.samples, self.reasons = index, [], set()
        self.max_mhz, self._stop, self._thr, self._record = None, threading.Event(), None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            if self._record:
                try:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for b, name in self.REASONS.items():
                        if bits & b:
                            self.reasons.add(name)
                except Exception:  # noqa: BLE001
                    pass
            time.sleep(0.005)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def begin(self):
        self._record = True

    def end(self):
        self._record = False

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=1.0)

    def summary(self):
        if self.nv is None:
            return self._smi_fallback()
        if not self.samples:  # timed region shorter than one period: take one sample now
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
            except Exception:  # noqa: BLE001
                pass
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}

    def _smi_fallback(self):
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=clocks.sm,clocks.max.sm",
                                  "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
            a, b = [int(x) for x in out.strip().split(",")]
            return {"sm_mhz": a, "sm_max_mhz": b, "reasons": [], "samples": 1}
        except Exception:  # noqa: BLE001
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}


def reference_arm(args) -> int:
    """The unmodified reference needs TensorFlow 1.x (tf.app.flags, tf.contrib, SyncReplicasOptimizer); it is a
    bare script without packaging, so `pip install /root/reference` has nothing to build (DESIGN.md records the
    attempt) and no TensorFlow wheel exists offline."""
    why = None
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    try:
        import tensorflow  # noqa: F401
    except Exception as exc:  # noqa: BLE001
        why = f"TensorFlow 1.x is not installable offline ({type(exc).__name__}: {exc}); reference is a bare TF1 script with no setup.py/pyproject, baseline/_ref holds no package"
    if why is None and not os.path.isdir(ref_dir):
        why = "baseline/_ref is absent: /root/reference has no setup.py/pyproject, pip cannot install it"
    if why is None:
        why = "reference script needs tf.contrib / tf.train.SyncReplicasOptimizer (TF1-only APIs)"
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


# BASELINE.md §3 (operative baseline: stock PyTorch + cuBLAS + NCCL on the same box, measured in round 1; the reference
# itself publishes no number and cannot run here). examples/s of the strongest stock arm per GPU count: the whole-step
# CUDA graph of PyTorch ops on one GPU, DistributedDataParallel over NCCL beyond.
BASELINE_MD = {
    "mnist_sync": {1: 511042.0, 2: 191551.0, 4: 345526.0, 8: 677315.0},
    "mnist_async": {2: 191551.0, 4: 345526.0, 8: 677315.0},
}


def stock_arm(args, rank, world, local_rank):
    """The matching stock-PyTorch arm, measured in this same invocation (same box, same steps / warm-up): the
    whole-step CUDA graph of PyTorch/cuBLAS kernels on one GPU, DDP over NCCL on several. Returns the arm's JSON
    dict on rank 0 (None elsewhere / when there is no arm for the config)."""
    import contextlib
    import io
    buf = io.StringIO()
    sub = argparse.Namespace(**vars(args))
    sub.steps = min(args.steps, 400)  # the eager DDP arm runs ~1.5 ms per step
    try:
        with contextlib.redirect_stdout(buf):
            if world == 1:
                from baseline.zero1_mlp import run_torch_graph_bench
                run_torch_graph_bench(sub, rank, world, local_rank)
            else:
                from baseline.ddp_mlp import run_ddp_bench
                run_ddp_bench(sub, rank, world, local_rank, keep_group=True)
    except Exception as exc:  # noqa: BLE001 - the product's own number must survive a failing baseline arm
        return {"impl": "stock", "unavailable": f"{type(exc).__name__}: {exc}"} if rank == 0 else None
    if rank != 0:
        return None
    for line in buf.getvalue().splitlines():
        if line.startswith("{"):
            try:
                return json.loads(line)
            except ValueError:
                pass
    return None


def replicas_identical(trainer, world) -> bool:
    """Every replica must hold bit-identical compute copies (what the next forward reads) and the same global_step."""
    import torch
    import torch.distributed as dist
    st = trainer.store
    parts = [t.view(torch.int16).to(torch.int64).sum() for t in st.w_t if t is not None]
    parts += [t.view(torch.int16).to(torch.int64).sum() for t in st.w_rm if t is not None]
    parts += [b.view(torch.int32).to(torch.int64).sum() for b in st.bias_c]
    parts.append(st.state[0].to(torch.int64))
    digest = torch.stack(parts)
    if world == 1:
        return True
    all_d = [torch.zeros_like(digest) for _ in range(world)]
    dist.all_gather(all_d, digest)
    return all(torch.equal(all_d[0], d) for d in all_d)


def run(args) -> int:
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)

    if args.impl == "ddp":
        from baseline.ddp_mlp import run_ddp_bench
        return run_ddp_bench(args, rank, world, local_rank)
    if args.impl == "zero1":
        from baseline.zero1_mlp import run_zero1_bench
        return run_zero1_bench(args, rank, world, local_rank)
    if args.impl == "torch_graph":
        from baseline.zero1_mlp import run_torch_graph_bench
        return run_torch_graph_bench(args, rank, world, local_rank)

    from tensorflow_train_distributed_b200.config import TrainConfig
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer

    comm = None
    if world > 1:
        from tensorflow_train_distributed_b200.parallel.symmetric import SymmetricComm
        comm = SymmetricComm.from_env()

    spec = CONFIGS[args.config]
    hosts = ",".join(f"127.0.0.1:{2222 + i}" for i in range(world))
    cfg = TrainConfig(job_name="worker", task_id=rank, worker_hosts=hosts, ps_hosts="127.0.0.1:2221",
                      model=spec["model"], sync_replicas=spec["sync_replicas"], batch_size=args.batch,
                      training_steps=10**9, seed=1234, save_checkpoint_secs=None, cuda_graph=not args.no_graph,
                      pdl=not args.no_pdl, fused_tail=args.fused_tail, pull_fused=args.pull_fused and not args.no_pull_fused,
                      step_impl=args.step_impl)
    trainer = GpuTrainer(cfg, device_index=local_rank, comm=comm)
    trainer.initialize()
    B, F = cfg.batch_size, cfg.input_dim

    # ---- synthetic data: a device-resident pool larger than L2 (126 MB) for the device-timed number, and a
    # pinned host pool for the end-to-end number ------------------------------------------------------------
    pool_batches = max(2, -(-160 * 2**20 // (B * F * 4)))
    ds = synthetic_mnist(num_examples=min(55000, pool_batches * B), input_dim=F, seed=100 + rank)
    reps = -(-pool_batches * B // ds.num_examples)
    xs = torch.from_numpy(ds.images).repeat(reps, 1)[:pool_batches * B].view(pool_batches, B, F).contiguous()
    ys = torch.from_numpy(ds.labels).repeat(reps)[:pool_batches * B].view(pool_batches, B).contiguous()
    host_pool = min(pool_batches, 64)
    hx = xs[:host_pool].clone().pin_memory()
    hy = ys[:host_pool].clone().pin_memory()
    with torch.cuda.stream(trainer.stream):
        dx, dy = xs.to(device), ys.to(device)
    trainer.use_device_dataset(dx, dy)

    sampler = ClockSampler(local_rank)
    sampler.start()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        sampler.begin()
        with torch.cuda.stream(trainer.stream):
            e0.record()
            for i in range(steps):
                fn(i)
            e1.record()
        barrier()
        sampler.end()
        ms = e0.elapsed_time(e1)
        if world > 1:
            ms = comm.max_over_ranks(ms)
        return ms

    # ---- device-timed: dataset resident on the GPU, graph replay per step --------------------------------
    def dev_step(i):
        trainer.run_step()

    with torch.cuda.stream(trainer.stream):
        for i in range(args.warmup):
            dev_step(i)
    ms = timed(dev_step, args.steps)
    values = trainer.read_values()

    # ---- end to end through the public API: pinned host batch -> H2D -> step -> D2H of the step's result ---
    trainer.use_device_dataset(None, None)
    slots = []

    def e2e_step(i):
        j = i % host_pool
        slots.append(trainer.step_async(hx[j], hy[j]))  # H2D (copy stream) + step + 64-byte D2H, one native call

    with torch.cuda.stream(trainer.stream):
        for i in range(args.warmup):
            e2e_step(i)
    slots.clear()
    ms_e2e = timed(e2e_step, args.steps)
    last = trainer.read_slot(slots[-1])
    sampler.stop()
    trainer.synchronize()
    same = replicas_identical(trainer, world) if (cfg.sync_replicas or world == 1) else None

    ex_per_step = B * world
    value = ex_per_step * args.steps / (ms / 1e3)
    e2e_value = ex_per_step * args.steps / (ms_e2e / 1e3)
    if rank == 0:
        out = {
            "metric": "examples/sec",
            "value": value,
            "unit": "examples/s",
            "n_gpus": world,
            "steps": args.steps,
            "warmup": args.warmup,
            "ms_per_step": ms / args.steps,
            "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": (value / BASELINE_MD[args.config][world]) if world in BASELINE_MD.get(args.config, {}) else None,
            "vs_baseline_source": "BASELINE.md section 3: strongest stock PyTorch arm on this box measured in round 1 (whole-step "
                                  "CUDA graph of cuBLAS kernels at 1 GPU, DDP+NCCL at 2/4/8); the reference publishes no number",
            "dtype": "bf16",
            "data": "synthetic MNIST-shaped (28x28x1 float32, 10 classes), random-init weights",
            "impl": "ours",
            "replicas_identical": same,
            "config": {
                "model": {"mlp": "distribute_training.py MLP 784-200-100-50-25-10 ELU dropout0.01",
                          "wide_mlp": "wide MLP 784-8192-8192-10", "lenet": "LeNet-5"}[cfg.model],
                "global_batch": ex_per_step, "batch_per_gpu": B, "seq_len": None,
                "parallelism": f"dp{world} " + ("sync replicas" if cfg.sync_replicas else "async replicas")
                               + ", GPU-sharded parameter server (flat shards, NVLink symmetric memory)",
                "optimizer": "SGD lr=0.01*0.96^floor(gs/468), L2 0.01, EMA 0.99",
                "l2_policy": f"inputs larger than L2: {pool_batches * B * F * 4 / 2**20:.0f} MiB device-resident pool rotated per step",
                "cuda_graph": cfg.cuda_graph,
                "pull_fused_first_gemm": bool(cfg.pull_fused and cfg.sync_replicas and world > 1 and cfg.model != "lenet"),
            },
            "e2e": {"value": e2e_value, "unit": "examples/s", "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": B * F * 4 + B * 8, "d2h_bytes_per_step": 64,
                    "path": "GpuTrainer.step_async per step: cudaMemcpyAsync of the pinned batch (x fp32 + labels int64) "
                            "on a copy stream into a 2-slot device pool, graph launch, 64-byte step state written to "
                            "pinned host memory by the update kernel; separate timed loop from `value`"},
            "gpu_launches": trainer.launches_per_step * args.steps,
            "launches_per_step": trainer.launches_per_step,
            "clocks": sampler.summary(),
            "final": {"loss": values.loss, "accuracy": values.accuracy, "global_step": values.global_step,
                      "e2e_loss": last.loss, "e2e_global_step": last.global_step},
        }
    trainer.close()
    stock = None
    if args.with_baseline and args.config in ("mnist_sync", "mnist_async", "wide_sync"):
        stock = stock_arm(args, rank, world, local_rank)  # same process group, after our own timed loops
    if rank == 0:
        if stock is not None:
            if "value" in stock:
                out["stock_same_run"] = {"impl": stock.get("impl"), "value": stock["value"], "ms_per_step": stock.get("ms_per_step"),
                                         "e2e_value": (stock.get("e2e") or {}).get("value"), "steps": stock.get("steps"),
                                         "ratio": value / stock["value"],
                                         "e2e_ratio": (e2e_value / stock["e2e"]["value"]) if stock.get("e2e") else None}
            else:
                out["stock_same_run"] = stock
        print(json.dumps(out))
    if world > 1:
        dist.barrier(device_ids=[local_rank])
        dist.destroy_process_group()
    return 0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "ddp", "zero1", "torch_graph"])
    ap.add_argument("--config", default="mnist_sync", choices=sorted(CONFIGS))
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-pdl", action="store_true")
    ap.add_argument("--fused-tail", action="store_true", help="run the narrow layers as one chain-kernel launch")
    ap.add_argument("--pull-fused", action="store_true",
                    help="fetch W1 from its owners inside the first GEMM (K3) instead of the update kernel's broadcast")
    ap.add_argument("--no-pull-fused", action="store_true", help="(default) K3 off")
    ap.add_argument("--step-impl", default="auto", choices=["auto", "column", "layers"],
                    help="auto/column: the column kernel where the model fits; layers: one GEMM launch per layer")
    ap.add_argument("--with-baseline", dest="with_baseline", action="store_true", default=True,
                    help="(default) also time the matching stock-PyTorch arm in this invocation -> stock_same_run")
    ap.add_argument("--no-baseline", dest="with_baseline", action="store_false")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return reference_arm(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch ourselves under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 200), os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    return run(args)


if __name__ == "__main__":
    sys.exit(main())
