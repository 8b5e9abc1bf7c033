"""Isolate what bounds the weight-streaming GEMM: a single CTA's k-loop rate, and row-pitch (partition camping)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from tensorflow_train_distributed_b200.ops import _native as N, gemm


def run(M, Nn, K, ldb_pad=0, bn=64, iters=10, no_swap=True):
    a = (torch.randn(M, K, device="cuda") * 0.02).to(torch.bfloat16)
    bfull = torch.zeros(Nn, K + ldb_pad, device="cuda", dtype=torch.bfloat16)
    bfull[:, :K] = (torch.randn(Nn, K, device="cuda") * 0.02).to(torch.bfloat16)
    out = torch.zeros(M, ((Nn + 7) // 8) * 8, device="cuda", dtype=torch.bfloat16)
    d = gemm.make_desc(a, bfull, M, Nn, K, N.EPI_FWD, out=out, block_n=bn)
    for _ in range(3):
        gemm.gemm_group([d], no_swap=no_swap, no_cluster=True)
    flush = torch.empty(256 * 2**20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gemm.gemm_group([d], no_swap=no_swap, no_cluster=True); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1000)
    ts.sort()
    return ts[len(ts) // 2]


def warm(M, Nn, K, bn=64, iters=20):
    a = (torch.randn(M, K, device="cuda") * 0.02).to(torch.bfloat16)
    b = (torch.randn(Nn, K, device="cuda") * 0.02).to(torch.bfloat16)
    out = torch.zeros(M, Nn, device="cuda", dtype=torch.bfloat16)
    d = gemm.make_desc(a, b, M, Nn, K, N.EPI_FWD, out=out, block_n=bn)
    for _ in range(3):
        gemm.gemm_group([d], no_swap=True, no_cluster=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        gemm.gemm_group([d], no_swap=True, no_cluster=True)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / iters


res = {"stages": os.environ.get("TTD_GEMM_STAGES"), "kspan": os.environ.get("TTD_GEMM_KSPAN")}
res["1cta_cold"] = run(128, 64, 8192)
res["1cta_warmL2"] = warm(128, 64, 8192)
res["128cta_cold"] = run(128, 8192, 8192)
print("STREAM " + json.dumps(res))
