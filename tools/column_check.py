"""Column kernel + fused-update net kernel against the per-layer path (1 GPU).

Two parameter stores with the same seed (identical weights), the same batches:
  (a) one forward/backward: activations, pre-activation gradients, weight gradients, loss / accuracy;
  (b) N full steps: master weights, shadows, compute copies, step state.
Prints one line per compared tensor and a JSON verdict; exits non-zero on mismatch.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from tensorflow_train_distributed_b200.models import mlp as mlp_model  # noqa: E402
from tensorflow_train_distributed_b200.ops import _native as N  # noqa: E402
from tensorflow_train_distributed_b200.parallel.param_store import GpuParamStore, Hyper  # noqa: E402


def build(sizes, B, column, drop, act="elu", pool=None, seed=11, ema=0.99):
    dev = torch.device("cuda", 0)
    layout = mlp_model.build_layout(sizes[0], sizes[1:])
    store = GpuParamStore(layout, Hyper(ema_decay=ema), dev)
    store.init(seed)
    model = mlp_model.GpuMlp(store, B, activation=act, dropout_rate=drop, seed=5, pdl=False, column=column, chain=False)
    if pool is not None:
        model.set_device_dataset(pool[0], pool[1])
    return store, model


def rel(a, b):
    a, b = a.float(), b.float()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-12))


def main() -> int:
    sizes = [int(x) for x in os.environ.get("SIZES", "784,200,100,50,25,10").split(",")]
    B = int(os.environ.get("BATCH", "128"))
    drop = float(os.environ.get("DROP", "0.1"))
    steps = int(os.environ.get("STEPS", "12"))
    P = 3
    torch.manual_seed(0)
    xs = torch.rand(P, B, sizes[0], device="cuda")
    ys = torch.randint(0, sizes[-1], (P, B), device="cuda")
    ok = True
    report = {}

    # ---- (a) one forward / backward ---------------------------------------------------------------------
    sa, ma = build(sizes, B, True, drop, pool=(xs, ys))
    sb, mb = build(sizes, B, False, drop, pool=(xs, ys))
    assert ma.use_column and not mb.use_column, (ma.use_column, mb.use_column)
    assert torch.equal(sa.w, sb.w)
    for m in (ma, mb):
        m.launch_forward_backward()
    torch.cuda.synchronize()
    L = ma.L
    for l in range(1, L):
        e = rel(ma.aT[l][:ma.sizes[l], :B], mb.aT[l][:mb.sizes[l], :B])
        print(f"aT[{l}]   rel err {e:.3e}")
        ok &= e < 2e-2
    for l in range(L):
        e = rel(ma.dzT[l][:, :B], mb.dzT[l][:, :B])
        print(f"dzT[{l}]  rel err {e:.3e}")
        ok &= e < 3e-2
    for i in range(L):
        e = rel(sa.grad_matrix(i), sb.grad_matrix(i))
        print(f"grad[{i}] rel err {e:.3e}")
        ok &= e < 3e-2
    ma_m = sa.state.view(torch.float32)[4:6].tolist()
    mb_m = sb.state.view(torch.float32)[4:6].tolist()
    print("metrics (loss, acc): column", ma_m, "per-layer", mb_m)
    ok &= abs(ma_m[0] - mb_m[0]) < 2e-3 * max(1.0, abs(mb_m[0])) and abs(ma_m[1] - mb_m[1]) < 1.5 / B
    report["fwd_bwd_ok"] = bool(ok)

    # ---- (b) full steps -------------------------------------------------------------------------------------
    sa, ma = build(sizes, B, True, drop, pool=(xs, ys))
    sb, mb = build(sizes, B, False, drop, pool=(xs, ys))
    for s, m in ((sa, ma), (sb, mb)):
        for _ in range(steps):
            m.launch_step()
    torch.cuda.synchronize()
    ew, ee = rel(sa.w, sb.w), rel(sa.ema, sb.ema)
    print(f"after {steps} steps: w rel err {ew:.3e}, ema rel err {ee:.3e}")
    ok2 = ew < 2e-3 and ee < 2e-3
    for i in range(L):
        e = rel(sa.w_t[i], sb.w_t[i])
        e2 = rel(sa.w_rm[i], sb.w_rm[i]) if sa.w_rm[i] is not None else 0.0
        e3 = rel(sa.bias_c[i], sb.bias_c[i]) if float(sb.bias_c[i].abs().max()) > 0 else 0.0
        print(f"copies[{i}]: w_t {e:.3e} w_rm {e2:.3e} bias {e3:.3e}")
        ok2 &= e < 1e-2 and e2 < 1e-2 and e3 < 1e-2
    sta = N.StepState.from_buffer_copy(sa.state.cpu().numpy().tobytes())
    stb = N.StepState.from_buffer_copy(sb.state.cpu().numpy().tobytes())
    print("state column   :", sta.global_step, sta.local_step, list(sta.metrics_out), sta.lr_last)
    print("state per-layer:", stb.global_step, stb.local_step, list(stb.metrics_out), stb.lr_last)
    ok2 &= sta.global_step == stb.global_step == steps and sta.local_step == stb.local_step == steps
    ok2 &= abs(sta.metrics_out[0] - stb.metrics_out[0]) < 5e-3 * max(1.0, stb.metrics_out[0]) and abs(sta.lr_last - stb.lr_last) < 1e-9
    report["steps_ok"] = bool(ok2)
    report["launches"] = "column: 2 per step; per-layer: 12"
    report["ok"] = bool(ok and ok2)
    print("COLUMN " + json.dumps(report))
    return 0 if report["ok"] else 1


if __name__ == "__main__":
    sys.exit(main())
