"""First-contact GPU probe: runs each kernel case in its own subprocess (a trapped kernel poisons its CUDA
context, so isolation keeps the other cases informative) and writes a report to gpurun_out/probe.json.

    python tools/gpu_probe.py            # all cases
    python tools/gpu_probe.py gemm_basic # one case, in-process
"""
import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
OUT = ROOT / "gpurun_out"
OUT.mkdir(exist_ok=True)


def _mk(rows, cols, ld=None, scale=1.0, seed=0):
    import torch
    ld = ld or ((cols + 7) // 8) * 8
    g = torch.Generator(device="cuda").manual_seed(seed)
    buf = torch.zeros(rows, ld, device="cuda", dtype=torch.bfloat16)
    buf[:, :cols] = (torch.randn(rows, cols, device="cuda", generator=g) * scale).to(torch.bfloat16)
    return buf


def gemm_case(M, Nn, K, block_n=0, dump=None):
    import torch
    from tensorflow_train_distributed_b200.ops import _native as N, gemm
    a = _mk(M, K, seed=1)
    b = _mk(Nn, K, seed=2)
    ldf = ((Nn + 3) // 4) * 4
    outf = torch.zeros(M, ldf, device="cuda")
    d = gemm.make_desc(a, b, M, Nn, K, N.EPI_WGRAD, outf=outf, scale=1.0, block_n=block_n)
    gemm.gemm_group([d])
    torch.cuda.synchronize()
    ref = a[:, :K].float() @ b[:, :K].float().t()
    err = (outf[:, :Nn] - ref).abs().max().item()
    tol = 1e-2 * max(1.0, ref.abs().max().item())
    ok = err <= tol
    if not ok and dump:
        torch.save({"a": a.cpu(), "b": b.cpu(), "out": outf.cpu(), "ref": ref.cpu()}, OUT / f"{dump}.pt")
    return ok, {"max_err": err, "tol": tol}


def case_gemm_basic():
    return gemm_case(128, 128, 64, dump="gemm_basic")


def case_gemm_k256():
    return gemm_case(128, 128, 256, dump="gemm_k256")


def case_gemm_n16():
    return gemm_case(128, 16, 64, dump="gemm_n16")


def case_gemm_n32_k25():
    return gemm_case(128, 25, 50, dump="gemm_n32")


def case_gemm_n256():
    return gemm_case(128, 256, 128, block_n=256, dump="gemm_n256")


def case_gemm_ragged():
    return gemm_case(785, 200, 128, dump="gemm_ragged")


def case_gemm_k784():
    return gemm_case(128, 200, 784, dump="gemm_k784")


def case_gemm_many_tiles():
    return gemm_case(512, 1024, 2048, dump=None)


def case_gemm_long_k():
    return gemm_case(128, 512, 8192, dump=None)


def case_fwd_elu():
    import torch
    from tensorflow_train_distributed_b200.ops import _native as N, gemm
    M, Nn, K = 128, 200, 784
    a, b = _mk(M, K, scale=0.05, seed=3), _mk(Nn, K, seed=4)
    bias = torch.randn(Nn, device="cuda")
    out = torch.zeros(M, 200, device="cuda", dtype=torch.bfloat16)
    outT = torch.zeros(Nn + 1, M, device="cuda", dtype=torch.bfloat16)
    d = gemm.make_desc(a, b, M, Nn, K, N.EPI_FWD, act=N.ACT_ELU, bias=bias, out=out, outT=outT)
    gemm.gemm_group([d])
    torch.cuda.synchronize()
    ref = torch.nn.functional.elu(a[:, :K].float() @ b[:, :K].float().t() + bias)
    e1 = (out.float() - ref).abs().max().item()
    e2 = (outT[:Nn].float().t() - ref).abs().max().item()
    tol = 2e-2 * max(1.0, ref.abs().max().item())
    return e1 <= tol and e2 <= tol, {"err_out": e1, "err_outT": e2, "tol": tol}


def case_dropout_dgrad():
    """Forward with dropout, then a dgrad whose mask must match the forward one."""
    import torch
    from tensorflow_train_distributed_b200.ops import _native as N, gemm
    M, Nn, K, rate = 128, 200, 256, 0.25
    a, b = _mk(M, K, scale=0.1, seed=5), _mk(Nn, K, seed=6)
    out = torch.zeros(M, 200, device="cuda", dtype=torch.bfloat16)
    act = torch.zeros(M, 200, device="cuda")
    step = torch.tensor([7], device="cuda", dtype=torch.int64)
    d = gemm.make_desc(a, b, M, Nn, K, N.EPI_FWD, act=N.ACT_ELU, drop_rate=rate, layer_id=3, out=out, outf=act)
    gemm.gemm_group([d], seed=1234, step_ptr=step.data_ptr())
    torch.cuda.synchronize()
    ref = torch.nn.functional.elu(a.float() @ b.float().t())
    kept = out.float() != 0
    frac = 1.0 - kept.float().mean().item()
    e_act = (act - ref).abs().max().item()
    e_keep = ((out.float() - ref / (1 - rate)) * kept).abs().max().item()
    # dgrad: dA_prev[M, Nn] = dZ[M, K2] . W[Nn, K2]^T, masked by the same dropout pattern
    K2 = 64
    dz, w = _mk(M, K2, seed=7), _mk(Nn, K2, seed=8)
    dprev = torch.zeros(M, 200, device="cuda", dtype=torch.bfloat16)
    d2 = gemm.make_desc(dz, w, M, Nn, K2, N.EPI_DGRAD, act=N.ACT_ELU, drop_rate=rate, layer_id=3, out=dprev, aux=out)
    gemm.gemm_group([d2], seed=1234, step_ptr=step.data_ptr())
    torch.cuda.synchronize()
    a_und = out.float() * (1 - rate)
    gref = (dz.float() @ w.float().t()) / (1 - rate) * torch.where(a_und > 0, torch.ones_like(a_und), a_und + 1) * kept
    e_dg = (dprev.float() - gref).abs().max().item()
    tol = 3e-2 * max(1.0, gref.abs().max().item())
    ok = abs(frac - rate) < 0.02 and e_act < 2e-2 * max(1, ref.abs().max().item()) and e_keep < 5e-2 * max(1, ref.abs().max().item()) and e_dg < tol
    return ok, {"drop_frac": frac, "e_act": e_act, "e_keep": e_keep, "e_dgrad": e_dg, "tol": tol}


def case_xent():
    import torch
    from tensorflow_train_distributed_b200.ops import _native as N, gemm
    M, Nn, K = 128, 10, 25
    a, b = _mk(M, K, ld=32, seed=9), _mk(Nn, K, ld=32, seed=10)
    bias = torch.randn(Nn, device="cuda")
    labels = torch.randint(0, Nn, (M,), device="cuda")
    dl = torch.zeros(M, 16, device="cuda", dtype=torch.bfloat16)
    dlT = torch.zeros(Nn + 1, M, device="cuda", dtype=torch.bfloat16)
    logits = torch.zeros(M, 16, device="cuda")
    metrics = torch.zeros(2, device="cuda")
    d = gemm.make_desc(a, b, M, Nn, K, N.EPI_FWD_XENT, bias=bias, out=dl, outT=dlT, outf=logits, labels=labels,
                       metrics_ptr=metrics.data_ptr())
    gemm.gemm_group([d])
    torch.cuda.synchronize()
    z = (a[:, :K].float() @ b[:, :K].float().t() + bias).requires_grad_(True)
    loss = torch.nn.functional.cross_entropy(z, labels)
    loss.backward()
    acc = (z.argmax(1) == labels).float().mean().item()
    e_l = abs(metrics[0].item() - loss.item())
    e_a = abs(metrics[1].item() - acc)
    e_g = (dl[:, :Nn].float() - z.grad).abs().max().item()
    e_gT = (dlT[:Nn].float().t() - z.grad).abs().max().item()
    e_z = (logits[:, :Nn] - z.detach()).abs().max().item()
    ok = e_l < 1e-3 and e_a < 1e-6 and e_g < 1e-4 and e_gT < 1e-4 and e_z < 1e-2
    # stand-alone kernel on the same logits
    m2 = torch.zeros(2, device="cuda")
    dl2 = torch.zeros(M, 16, device="cuda", dtype=torch.bfloat16)
    gemm.softmax_xent(logits, labels, Nn, dl2, None, m2)
    torch.cuda.synchronize()
    e_l2 = abs(m2[0].item() - loss.item())
    ok = ok and e_l2 < 1e-2 and (dl2[:, :Nn].float() - z.grad).abs().max().item() < 1e-4
    return ok, {"e_loss": e_l, "e_acc": e_a, "e_grad": e_g, "e_gradT": e_gT, "e_logits": e_z, "e_loss_standalone": e_l2}


def case_group2():
    """dgrad + wgrad of one layer in a single launch."""
    import torch
    from tensorflow_train_distributed_b200.ops import _native as N, gemm
    B, fin, fout = 128, 200, 100
    dz, dzT = _mk(B, fout, ld=104, seed=11), None
    dzT = torch.zeros(fout, B, device="cuda", dtype=torch.bfloat16)
    dzT[:, :] = dz[:, :fout].t()
    w = _mk(fin, fout, ld=104, seed=12)           # TF layout [in, out]
    aT = torch.ones(fin + 1, B, device="cuda", dtype=torch.bfloat16)
    aT[:fin] = _mk(fin, B, seed=13)
    dprev = torch.zeros(B, 200, device="cuda", dtype=torch.bfloat16)
    dw = torch.zeros(fin + 1, fout, device="cuda")
    d1 = gemm.make_desc(dz, w, B, fin, fout, N.EPI_DGRAD, out=dprev)
    d2 = gemm.make_desc(aT, dzT, fin + 1, fout, B, N.EPI_WGRAD, outf=dw)
    gemm.gemm_group([d1, d2])
    torch.cuda.synchronize()
    r1 = dz[:, :fout].float() @ w[:, :fout].float().t()
    r2 = aT.float() @ dzT.float().t()
    e1 = (dprev.float() - r1).abs().max().item()
    e2 = (dw - r2).abs().max().item()
    return e1 < 3e-2 * r1.abs().max().item() and e2 < 1e-2 * r2.abs().max().item(), {"e_dgrad": e1, "e_wgrad": e2}


def case_params():
    import ctypes as C
    import torch
    from tensorflow_train_distributed_b200.ops import _native as N
    lib = N.load()
    shapes = [(784, 200), (200, 100), (100, 50), (50, 25), (25, 10)]
    off, vars_, copies = 0, [], []
    for i, (r, c) in enumerate(shapes):
        v = N.VarDesc()
        v.offset, v.rows, v.cols, v.regularize = off, r, c, 1 if i < 4 else 0
        ldr, ldt = ((c + 7) // 8) * 8, ((r + 7) // 8) * 8
        wrm = torch.zeros(r + 1, ldr, device="cuda", dtype=torch.bfloat16)
        wt = torch.zeros(c, ldt, device="cuda", dtype=torch.bfloat16)
        v.w_rm, v.ld_rm, v.w_t, v.ld_t = wrm.data_ptr(), ldr, wt.data_ptr(), ldt
        copies.append((wrm, wt))
        vars_.append(v)
        off += ((r + 1) * c + 63) // 64 * 64
    total = off
    arr = (N.VarDesc * len(vars_))(*vars_)
    w = torch.zeros(total, device="cuda")
    ema = torch.zeros(total, device="cuda")
    N.check(lib.ttd_init_params(w.data_ptr(), ema.data_ptr(), arr, len(vars_), 42, N.cur_stream()), "init")
    torch.cuda.synchronize()
    k1 = w[:784 * 200]
    std, expect = k1.std().item(), (2.6 / 784) ** 0.5 * 0.8796  # std of a unit normal truncated at 2 sigma
    info = {"std": std, "expect": expect, "absmax_over_sigma": k1.abs().max().item() / (2.6 / 784) ** 0.5,
            "bias_absmax": w[784 * 200:784 * 200 + 200].abs().max().item()}
    ok = abs(std - expect) / expect < 0.03 and info["absmax_over_sigma"] <= 2.0001 and info["bias_absmax"] == 0
    ok = ok and torch.equal(w, ema)
    wrm, wt = copies[0]
    ok = ok and torch.equal(wrm[:784, :200].float(), k1.view(784, 200).to(torch.bfloat16).float())
    ok = ok and torch.equal(wt[:, :784].float(), k1.view(784, 200).t().to(torch.bfloat16).float())
    # one update step against a torch oracle
    state = torch.zeros(8, device="cuda", dtype=torch.int64)
    state[0] = 500  # global_step -> one decay
    g = torch.randn(total, device="cuda") * 0.1
    h = N.Hyper(0.01, 0.96, 468, 1, 2, 0.01, 0.99, 1, 1.0)
    p = N.Peers()
    p.grads[0] = g.data_ptr()
    p.n_grads = p.n_copies = 1
    w0, ema0 = w.clone(), ema.clone()
    N.check(lib.ttd_apply_update(w.data_ptr(), ema.data_ptr(), arr, len(vars_), 0, total, C.byref(h), C.byref(p),
                                 state.data_ptr(), 1, N.cur_stream()), "update")
    torch.cuda.synchronize()
    lr = 0.01 * 0.96 ** (500 // 468)
    regmask = torch.zeros(total, device="cuda")
    for v in vars_[:4]:
        regmask[v.offset:v.offset + v.rows * v.cols] = 1
    valid = torch.zeros(total, device="cuda", dtype=torch.bool)
    for v in vars_:
        valid[v.offset:v.offset + (v.rows + 1) * v.cols] = True
    wref = w0 - lr * (g + 0.01 * w0 * regmask)
    dec = min(0.99, (1 + 501) / (10 + 501))
    eref = ema0 - (1 - dec) * (ema0 - wref)
    e_w = ((w - wref) * valid).abs().max().item()
    e_e = ((ema - eref) * valid).abs().max().item()
    st = state.cpu()
    info.update({"e_w": e_w, "e_ema": e_e, "gs": int(st[0]), "ls": int(st[1])})
    ok = ok and e_w < 1e-6 and e_e < 1e-6 and int(st[0]) == 501 and int(st[1]) == 1
    ok = ok and torch.equal(wt[:, :784].float(), w[:784 * 200].view(784, 200).t().to(torch.bfloat16).float())
    return ok, info


def case_bench_wide():
    """Timing of the weight-streaming GEMM of the wide MLP: [128, 8192] x [8192, 8192]."""
    import torch
    from tensorflow_train_distributed_b200.ops import _native as N, gemm
    M, Nn, K = 128, 8192, 8192
    a, b = _mk(M, K, scale=0.02, seed=1), _mk(Nn, K, scale=0.02, seed=2)
    out = torch.zeros(M, Nn, device="cuda", dtype=torch.bfloat16)
    info = {}
    for bn in (32, 64, 128):
        d = gemm.make_desc(a, b, M, Nn, K, N.EPI_FWD, act=N.ACT_RELU, out=out, block_n=bn)
        for _ in range(3):
            gemm.gemm_group([d])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            gemm.gemm_group([d])
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1000 / 20
        info[f"bn{bn}_us"] = us
        info[f"bn{bn}_GBs"] = (Nn * K * 2 + M * K * 2 + M * Nn * 2) / us / 1e3
        info[f"bn{bn}_TF"] = 2.0 * M * Nn * K / us / 1e6
    ref = torch.relu(a.float() @ b.float().t())
    err = (out.float() - ref).abs().max().item()
    info["err"] = err
    t0 = time.time()
    for _ in range(3):
        torch.matmul(a, b.t())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        torch.matmul(a, b.t())
    e1.record()
    torch.cuda.synchronize()
    info["cublas_us"] = e0.elapsed_time(e1) * 1000 / 20
    return err < 0.05 * max(1.0, ref.abs().max().item()), info


CASES = {k[5:]: v for k, v in list(globals().items()) if k.startswith("case_")}


def main():
    if len(sys.argv) > 1 and sys.argv[1] != "--all":
        name = sys.argv[1]
        try:
            ok, info = CASES[name]()
        except Exception as exc:  # noqa: BLE001
            ok, info = False, {"exception": repr(exc)[:500]}
        print("RESULT " + json.dumps({"case": name, "ok": bool(ok), **info}))
        return 0 if ok else 1
    report = []
    for name in CASES:
        t0 = time.time()
        try:
            p = subprocess.run([sys.executable, __file__, name], capture_output=True, text=True, timeout=180)
            line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")]
            rec = json.loads(line[-1][7:]) if line else {"case": name, "ok": False, "stderr": p.stderr[-800:], "rc": p.returncode}
        except subprocess.TimeoutExpired:
            rec = {"case": name, "ok": False, "timeout": True}
        rec["secs"] = round(time.time() - t0, 1)
        report.append(rec)
        print(json.dumps(rec), flush=True)
    (OUT / "probe.json").write_text(json.dumps(report, indent=1))
    return 0 if all(r["ok"] for r in report) else 1


if __name__ == "__main__":
    sys.exit(main())
