"""Stage timeline of the two-launch step (column kernel + weight-gradient/update net kernel), 1 GPU.

globaltimer stamps written by CTA 0 of the column kernel (dependency resolved, accumulator of stage t ready, stage
t stored, end) and by the net kernel (start after its dependency, per problem: operands landed / epilogue done /
counted in, last CTA done). Stamps of the LAST replay of a CUDA-graph loop are shown, so the numbers include the
programmatic-dependent-launch overlap the bench sees. Prints microseconds relative to the column kernel's start.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from tensorflow_train_distributed_b200.config import TrainConfig  # noqa: E402
from tensorflow_train_distributed_b200.data import synthetic_mnist  # noqa: E402
from tensorflow_train_distributed_b200.ops import _native as N  # noqa: E402
from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer  # noqa: E402


def main() -> None:
    lib = N.load()
    ctrace = torch.zeros(64, dtype=torch.int64, device="cuda")
    ntrace = torch.zeros(4 * 24 + 2, dtype=torch.int64, device="cuda")
    lib.ttd_mlp_column_set_trace(ctrace.data_ptr())   # captured into the graph's kernel parameters
    lib.ttd_gemm_net_set_trace(ntrace.data_ptr())
    cfg = TrainConfig(worker_hosts="127.0.0.1:2222", save_checkpoint_secs=None, seed=3)
    tr = GpuTrainer(cfg)
    tr.initialize()
    ds = synthetic_mnist(128 * 8, seed=4)
    xs = torch.tensor(ds.images).view(8, 128, 784).cuda()
    ys = torch.tensor(ds.labels).view(8, 128).cuda()
    tr.use_device_dataset(xs, ys)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(tr.stream):
        for _ in range(20):
            tr.run_step()
        e0.record()
        for _ in range(500):
            tr.run_step()
        e1.record()
    tr.synchronize()
    us_step = e0.elapsed_time(e1) * 1e3 / 500
    c = ctrace.cpu().tolist()
    n = ntrace.cpu().tolist()
    t0 = c[0]
    L = len(cfg.layer_sizes)
    rel = lambda v: round((v - t0) / 1e3, 2) if v > 0 else None  # noqa: E731
    names = [f"F{l + 1}" for l in range(L)] + [f"D{l}" for l in range(L, 1, -1)]
    print(f"step (graph replay, with tracing on): {us_step:.2f} us")
    print("column kernel, CTA 0 (us after its dependency resolved): stage  accumulator-ready  stored")
    for t, nm in enumerate(names):
        print(f"  {nm:>3}  {rel(c[1 + 2 * t])}  {rel(c[2 + 2 * t])}")
    print("  end", rel(c[63]))
    print("net kernel: start after dependency", rel(n[96]), " last CTA done", rel(n[97]))
    wn = [f"W{l}" for l in range(L, 0, -1)]
    for i, nm in enumerate(wn):
        row = n[4 * i:4 * i + 4]
        print(f"  {nm}: deps met {rel(row[0])} operands landed {rel(row[1])} epilogue done {rel(row[2])} counted in {rel(row[3])}")
    print("TRACE " + json.dumps({"us_per_step": us_step, "column": [rel(v) for v in c[:2 * len(names) + 1]] + [rel(c[63])],
                                 "net_start": rel(n[96]), "net_done": rel(n[97])}))
    lib.ttd_mlp_column_set_trace(None)
    lib.ttd_gemm_net_set_trace(None)
    tr.close()


if __name__ == "__main__":
    main()
