#!/usr/bin/env python
"""One training step of each model inside a cudaProfilerStart/Stop window, for
`ncu --profile-from-start off --set full ... python tools/profile_step.py` (single GPU, no CUDA graph)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from tensorflow_train_distributed_b200.config import TrainConfig  # noqa: E402
from tensorflow_train_distributed_b200.data import synthetic_mnist  # noqa: E402
from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer  # noqa: E402


def main() -> None:
    models = os.environ.get("PROFILE_MODELS", "mlp,lenet,wide_mlp").split(",")
    ds = synthetic_mnist(1024, seed=3)
    for m in models:
        cfg = TrainConfig(model=m, sync_replicas=True, worker_hosts="127.0.0.1:2222", save_checkpoint_secs=None,
                          cuda_graph=False, activation="relu" if m == "wide_mlp" else "elu")
        tr = GpuTrainer(cfg)
        tr.initialize()
        for _ in range(3):
            xs, ys = ds.next_batch(cfg.batch_size)
            tr.train_step(xs, ys)
        xs, ys = ds.next_batch(cfg.batch_size)
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        v = tr.train_step(xs, ys)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        print(f"PROFILED {m}: loss {v.loss:.4f} global_step {v.global_step} kernels/step {tr.launches_per_step}", flush=True)
        tr.close()


if __name__ == "__main__":
    main()
