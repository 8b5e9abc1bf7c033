"""Which part of the FWD/DGRAD epilogue is slow? Time a one-tile GEMM (128x112x64) with features toggled."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from tensorflow_train_distributed_b200.ops import _native as N, gemm

M, Nn, K = 128, 112, 64
a = (torch.randn(M, K, device="cuda") * 0.1).to(torch.bfloat16)
b = (torch.randn(Nn, K, device="cuda") * 0.1).to(torch.bfloat16)
bias = torch.randn(Nn, device="cuda")
out = torch.zeros(M, Nn, device="cuda", dtype=torch.bfloat16)
outT = torch.zeros(Nn + 1, M, device="cuda", dtype=torch.bfloat16)
outf = torch.zeros(M, Nn, device="cuda")
aux = (torch.randn(M, Nn, device="cuda")).to(torch.bfloat16)
step = torch.tensor([3], device="cuda", dtype=torch.int64)


def t(desc, iters=300):
    g = torch.cuda.CUDAGraph()
    for _ in range(3):
        gemm.gemm_group([desc], seed=1, step_ptr=step.data_ptr())
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(50):
            gemm.gemm_group([desc], seed=1, step_ptr=step.data_ptr())
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters // 50):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) * 1000 / (iters // 50 * 50), 2)


mk = gemm.make_desc
res = {}
res["wgrad_fp32"] = t(mk(a, b, M, Nn, K, N.EPI_WGRAD, outf=outf))
res["fwd_none_out_only"] = t(mk(a, b, M, Nn, K, N.EPI_FWD, out=out))
res["fwd_none_outT_only"] = t(mk(a, b, M, Nn, K, N.EPI_FWD, outT=outT))
res["fwd_none_out_outT"] = t(mk(a, b, M, Nn, K, N.EPI_FWD, out=out, outT=outT))
res["fwd_bias"] = t(mk(a, b, M, Nn, K, N.EPI_FWD, bias=bias, out=out, outT=outT))
res["fwd_bias_elu"] = t(mk(a, b, M, Nn, K, N.EPI_FWD, act=N.ACT_ELU, bias=bias, out=out, outT=outT))
res["fwd_bias_elu_drop"] = t(mk(a, b, M, Nn, K, N.EPI_FWD, act=N.ACT_ELU, drop_rate=0.01, layer_id=1, bias=bias, out=out, outT=outT))
res["fwd_drop_only"] = t(mk(a, b, M, Nn, K, N.EPI_FWD, drop_rate=0.01, layer_id=1, out=out))
res["dgrad_elu_drop"] = t(mk(a, b, M, Nn, K, N.EPI_DGRAD, act=N.ACT_ELU, drop_rate=0.01, layer_id=1, aux=aux, out=out, outT=outT))
res["dgrad_elu"] = t(mk(a, b, M, Nn, K, N.EPI_DGRAD, act=N.ACT_ELU, aux=aux, out=out, outT=outT))
print("EPI " + json.dumps(res))
