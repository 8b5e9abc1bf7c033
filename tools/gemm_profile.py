"""Small driver for ncu captures and timing of the tcgen05 GEMM on the wide-MLP shapes.

    python tools/gemm_profile.py [fwd|wgrad|dgrad] [iters] [block_n]
"""
import json
import sys
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from tensorflow_train_distributed_b200.ops import _native as N, gemm  # noqa: E402


def mk(rows, cols, seed, scale=0.02):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(rows, cols, device="cuda", generator=g) * scale).to(torch.bfloat16)


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "fwd"
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    bn = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    B, H = int(os.environ.get("GP_BATCH", "128")), 8192
    if which == "fwd":      # Z[B,H] = A[B,H] . Wt[H,H]^T  (+bias, relu)
        a, b = mk(B, H, 1), mk(H, H, 2)
        out = torch.zeros(B, H, device="cuda", dtype=torch.bfloat16)
        outT = torch.zeros(H + 1, B, device="cuda", dtype=torch.bfloat16)
        bias = torch.zeros(H, device="cuda")
        d = gemm.make_desc(a, b, B, H, H, N.EPI_FWD, act=N.ACT_RELU, bias=bias, out=out, outT=outT, block_n=bn)
        flops, bytes_ = 2.0 * B * H * H, H * H * 2 + B * H * 2 * 3
    elif which == "dgrad":
        a, b = mk(B, H, 1), mk(H, H, 2)
        aux = mk(B, H, 3).abs()
        out = torch.zeros(B, H, device="cuda", dtype=torch.bfloat16)
        outT = torch.zeros(H, B, device="cuda", dtype=torch.bfloat16)
        d = gemm.make_desc(a, b, B, H, H, N.EPI_DGRAD, act=N.ACT_RELU, aux=aux, out=out, outT=outT, block_n=bn)
        flops, bytes_ = 2.0 * B * H * H, H * H * 2 + B * H * 2 * 4
    else:                   # dW[H+1,H] = At[H+1,B] . dZt[H,B]^T  -> fp32
        a, b = mk(H + 1, B, 1), mk(H, B, 2)
        outf = torch.zeros(H + 1, H, device="cuda")
        d = gemm.make_desc(a, b, H + 1, H, B, N.EPI_WGRAD, outf=outf, block_n=bn)
        flops, bytes_ = 2.0 * B * H * H, H * H * 4 + 2 * B * H * 2
    for _ in range(3):
        gemm.gemm_group([d])
    torch.cuda.synchronize()
    flush = torch.empty(256 * 2**20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(iters):
        flush.zero_()  # evict L2 (126 MB) between timed launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        gemm.gemm_group([d])
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1000)
    ts.sort()
    us = ts[len(ts) // 2]
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
    print("GEMMPROF " + json.dumps({"which": which, "block_n": bn, "us_median": us, "us_min": ts[0], "TFLOPs": flops / us / 1e6,
                                    "GBs": bytes_ / us / 1e3, "frac_hbm_of_measured": bytes_ / us / 1e3 / peaks["hbm_gbs"],
                                    "frac_bf16_of_measured": flops / us / 1e6 / peaks["bf16_tflops"]}))


if __name__ == "__main__":
    main()
