#!/usr/bin/env python
"""Net kernel (all GEMMs of a step in one launch) against the per-layer path: equality, timing, stage trace."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from tensorflow_train_distributed_b200.config import TrainConfig  # noqa: E402
from tensorflow_train_distributed_b200.data import synthetic_mnist  # noqa: E402
from tensorflow_train_distributed_b200.ops import _native as N  # noqa: E402
from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer  # noqa: E402


def cfg(**kw):
    base = dict(worker_hosts="127.0.0.1:2222", batch_size=128, save_checkpoint_secs=None, seed=11, sync_replicas=True,
                model=os.environ.get("NET_MODEL", "mlp"))
    base.update(kw)
    return TrainConfig(**base)


def timed(tr, steps=1000):
    ds = synthetic_mnist(128 * 64, seed=1)
    xs = torch.from_numpy(ds.images[:128 * 64]).view(64, 128, -1).cuda()
    ys = torch.from_numpy(ds.labels[:128 * 64]).view(64, 128).cuda()
    tr.use_device_dataset(xs, ys)
    with torch.cuda.stream(tr.stream):
        for _ in range(50):
            tr.run_step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            tr.run_step()
        e1.record()
    tr.synchronize()
    return e0.elapsed_time(e1) * 1e3 / steps


def main():
    out = {}
    graph = os.environ.get("NET_GRAPH", "1") == "1"
    a = GpuTrainer(cfg(dropout_rate=0.1, fused_tail=True, tail_impl="net", cuda_graph=graph))
    b = GpuTrainer(cfg(dropout_rate=0.1, fused_tail=False, step_impl="layers", cuda_graph=graph))
    a.initialize(), b.initialize()
    out["launches"] = [a.launches_per_step, b.launches_per_step]
    out["net_ctas"] = getattr(a.model, "_net", None).ctas if hasattr(a.model, "_net") else a.model.head._net.ctas
    ds1, ds2 = synthetic_mnist(4096, seed=6), synthetic_mnist(4096, seed=6)
    for _ in range(30):
        va, vb = a.train_step(*ds1.next_batch(128)), b.train_step(*ds2.next_batch(128))
    out["loss"] = [va.loss, vb.loss]
    wa, wb = a.export_variables(), b.export_variables()
    out["max_param_diff"] = max(float(np.abs(wa[k] - wb[k]).max()) for k in wa if k != "global_step")
    out["ok"] = bool(abs(va.loss - vb.loss) < 1e-5 and out["max_param_diff"] < 1e-6)
    print("NETCHECK " + json.dumps(out), flush=True)
    # stage trace of one eager net launch
    net = a.model._net if hasattr(a.model, "_net") else a.model.head._net
    tbuf = torch.zeros(4 * net.n, dtype=torch.int64, device="cuda")
    N.load().ttd_gemm_net_set_trace(tbuf.data_ptr())
    t = GpuTrainer(cfg(dropout_rate=0.1, fused_tail=True, tail_impl="net", cuda_graph=False))
    t.initialize()
    for _ in range(3):
        t.train_step(*ds1.next_batch(128))
    tnet = t.model._net if hasattr(t.model, "_net") else t.model.head._net
    tbuf2 = torch.zeros(4 * tnet.n, dtype=torch.int64, device="cuda")
    N.load().ttd_gemm_net_set_trace(tbuf2.data_ptr())
    t.train_step(*ds1.next_batch(128))
    N.load().ttd_gemm_net_set_trace(None)
    tt = tbuf2.cpu().view(-1, 4)
    t0 = int(tt[tt > 0].min())
    print("NETTRACE us (deps met, operands landed, accumulator full+epilogue, counted in):",
          [[round((int(v) - t0) / 1e3, 2) if v > 0 else None for v in row] for row in tt.tolist()], flush=True)
    t.close()
    ta, tb = timed(a), timed(b)
    print("NETTIME " + json.dumps({"net_us_per_step": ta, "per_layer_us_per_step": tb}), flush=True)
    a.close(), b.close()


if __name__ == "__main__":
    main()
