#!/usr/bin/env python
"""Where the 1-GPU step goes: each part of the step captured as its own CUDA graph and replayed back to back
(device time per replay, warm caches, PDL inside each part as in the full step)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from tensorflow_train_distributed_b200.config import TrainConfig  # noqa: E402
from tensorflow_train_distributed_b200.data import synthetic_mnist  # noqa: E402
from tensorflow_train_distributed_b200.ops import gemm  # noqa: E402
from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer  # noqa: E402


def main() -> None:
    model = os.environ.get("BREAKDOWN_MODEL", "mlp")
    cfg = TrainConfig(model=model, sync_replicas=True, worker_hosts="127.0.0.1:2222", save_checkpoint_secs=None,
                      cuda_graph=False, activation="relu" if model == "wide_mlp" else "elu")
    tr = GpuTrainer(cfg)
    tr.initialize()
    ds = synthetic_mnist(1024, seed=3)
    for _ in range(3):
        tr.train_step(*ds.next_batch(cfg.batch_size))
    m, st = tr.model, tr.store
    head = getattr(m, "head", m)

    def part_prep():
        gemm.prep_input(head.x if hasattr(head, "x") else m.x, head.a[0], head.aT[0], pdl=cfg.pdl)

    parts = {
        "forward (all layers + xent)": lambda: head.launch_forward(),
        "backward (all layers)": lambda: head.launch_backward(),
        "update K1 (1 GPU)": lambda: st.apply_update(True, pdl=cfg.pdl),
        "fwd+bwd": lambda: (head.launch_forward(), head.launch_backward()),
        "full step": lambda: m.launch_step(sync=True),
    }
    if model != "lenet":
        parts["prep_input"] = part_prep
        for i, d in enumerate(head.fwd_descs):
            parts[f"fwd layer {i + 1} alone"] = (lambda d=d: gemm.gemm_group([d], seed=1, step_ptr=st.local_step_ptr, pdl=cfg.pdl))
        for i, grp in enumerate(head.bwd_descs):
            parts[f"bwd group {len(head.bwd_descs) - i} alone"] = (
                lambda grp=grp: gemm.gemm_group(grp, seed=1, step_ptr=st.local_step_ptr, pdl=cfg.pdl))
    out = {}

    def measure(fn) -> float:
        with torch.cuda.stream(tr.stream):
            fn()
        tr.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=tr.stream):
            fn()
        with torch.cuda.stream(tr.stream):
            for _ in range(20):
                g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(500):
                g.replay()
            e1.record()
        tr.synchronize()
        return round(e0.elapsed_time(e1) * 1e3 / 500, 2)

    for name, fn in parts.items():
        try:
            out[name] = measure(fn)
        except Exception as exc:  # noqa: BLE001 - keep the other parts
            out[name] = f"failed: {exc}"
    print("BREAKDOWN " + json.dumps({"model": model, "us_per_replay": out}), flush=True)
    tr.close()


if __name__ == "__main__":
    main()
