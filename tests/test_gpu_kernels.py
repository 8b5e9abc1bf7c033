"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference of the same op (GPU tier)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _mk(rows, cols, ld=None, scale=1.0, seed=0):
    ld = ld or ((cols + 7) // 8) * 8
    g = torch.Generator(device="cuda").manual_seed(seed)
    buf = torch.zeros(rows, ld, device="cuda", dtype=torch.bfloat16)
    buf[:, :cols] = (torch.randn(rows, cols, device="cuda", generator=g) * scale).to(torch.bfloat16)
    return buf


@pytest.fixture(scope="module")
def ops():
    from tensorflow_train_distributed_b200.ops import _native as N, gemm
    N.load()
    return N, gemm


@pytest.mark.parametrize("M,Nn,K,bn", [
    (128, 128, 64, 0), (128, 128, 256, 0), (128, 16, 64, 0), (128, 25, 50, 0), (128, 256, 128, 256),
    (785, 200, 128, 0), (128, 200, 784, 0), (201, 100, 128, 0), (512, 1024, 2048, 0), (128, 512, 8192, 64),
    (100, 10, 25, 0), (64, 8192, 784, 0),
    (128, 2048, 1024, 64),   # 32 N tiles -> clusters of 4 share the A tile by TMA multicast
    (100, 512, 512, 128),    # 4 N tiles -> clusters of 2
    (128, 8192, 8192, 64),   # the wide-MLP forward shape
    (300, 96, 128, 32), (1025, 512, 128, 0),  # wgrad-like: smem-staged, line-coalesced fp32 stores
])
def test_gemm_matches_fp32_matmul(ops, M, Nn, K, bn):
    N, gemm = ops
    a, b = _mk(M, K, seed=1), _mk(Nn, K, seed=2)
    outf = torch.zeros(M, ((Nn + 3) // 4) * 4, device="cuda")
    gemm.gemm_group([gemm.make_desc(a, b, M, Nn, K, N.EPI_WGRAD, outf=outf, scale=0.5, block_n=bn)])
    torch.cuda.synchronize()
    ref = 0.5 * (a[:, :K].float() @ b[:, :K].float().t())
    # inputs are exactly representable bf16 and accumulation is fp32: only summation order differs
    assert torch.allclose(outf[:, :Nn], ref, rtol=1e-4, atol=1e-3 * K ** 0.5)
    assert outf[:, Nn:].abs().max().item() == 0 if outf.shape[1] > Nn else True


@pytest.mark.parametrize("act", ["none", "relu", "elu"])
def test_forward_epilogue_bias_activation(ops, act):
    N, gemm = ops
    M, Nn, K = 128, 200, 784
    a, b = _mk(M, K, scale=0.05, seed=3), _mk(Nn, K, seed=4)
    bias = torch.randn(Nn, device="cuda")
    out = torch.zeros(M, 200, device="cuda", dtype=torch.bfloat16)
    outT = torch.zeros(Nn + 1, M, device="cuda", dtype=torch.bfloat16)
    outf = torch.zeros(M, 200, device="cuda")
    gemm.gemm_group([gemm.make_desc(a, b, M, Nn, K, N.EPI_FWD, act=N.ACT_CODES[act], bias=bias, out=out, outT=outT,
                                    outf=outf)])
    torch.cuda.synchronize()
    z = a[:, :K].float() @ b[:, :K].float().t() + bias
    ref = {"none": z, "relu": torch.relu(z), "elu": torch.nn.functional.elu(z)}[act]
    assert torch.allclose(outf, ref, rtol=1e-3, atol=1e-3)
    assert torch.allclose(out.float(), ref, rtol=1e-2, atol=1e-2)          # bf16 rounding of the stored copy
    assert torch.equal(outT[:Nn].t().contiguous(), out[:, :Nn].contiguous())  # both copies hold the same values
    assert outT[Nn].abs().max().item() == 0                                  # the ones row belongs to the caller


def test_dropout_statistics_and_dgrad_mask_agree(ops):
    N, gemm = ops
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
    assert abs((1.0 - kept.float().mean().item()) - rate) < 0.02
    assert torch.allclose(act, ref, rtol=1e-3, atol=1e-3)
    assert torch.allclose(out.float()[kept], (ref / (1 - rate))[kept], rtol=1e-2, atol=1e-2)
    # a different step gives a different mask, the same step the same mask
    out2 = torch.zeros_like(out)
    d2 = gemm.make_desc(a, b, M, Nn, K, N.EPI_FWD, act=N.ACT_ELU, drop_rate=rate, layer_id=3, out=out2)
    gemm.gemm_group([d2], seed=1234, step_ptr=step.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(out, out2)
    step += 1
    gemm.gemm_group([d2], seed=1234, step_ptr=step.data_ptr())
    torch.cuda.synchronize()
    assert not torch.equal(out, out2)
    step -= 1
    # dgrad regenerates the forward mask
    K2 = 64
    dz, w = _mk(M, K2, seed=7), _mk(Nn, K2, seed=8)
    dprev = torch.zeros(M, 200, device="cuda", dtype=torch.bfloat16)
    gemm.gemm_group([gemm.make_desc(dz, w, M, Nn, K2, N.EPI_DGRAD, act=N.ACT_ELU, drop_rate=rate, layer_id=3,
                                    out=dprev, aux=out)], seed=1234, step_ptr=step.data_ptr())
    torch.cuda.synchronize()
    a_und = out.float() * (1 - rate)
    gref = (dz.float() @ w.float().t()) / (1 - rate) * torch.where(a_und > 0, torch.ones_like(a_und), a_und + 1) * kept
    assert torch.allclose(dprev.float(), gref, rtol=2e-2, atol=2e-2)
    assert (dprev.float()[~kept] == 0).all()


@pytest.mark.parametrize("M,Nn,K", [(128, 10, 25), (100, 10, 84), (128, 100, 64)])
def test_fused_xent_matches_torch(ops, M, Nn, K):
    N, gemm = ops
    ldk = ((K + 7) // 8) * 8
    a, b = _mk(M, K, ld=ldk, seed=9), _mk(Nn, K, ld=ldk, seed=10)
    bias = torch.randn(Nn, device="cuda")
    labels = torch.randint(0, Nn, (M,), device="cuda")
    ldn = ((Nn + 7) // 8) * 8
    ldm = ((M + 7) // 8) * 8
    dl = torch.zeros(M, ldn, device="cuda", dtype=torch.bfloat16)
    dlT = torch.zeros(Nn, ldm, device="cuda", dtype=torch.bfloat16)
    logits = torch.zeros(M, ldn, device="cuda")
    metrics = torch.zeros(2, device="cuda")
    gemm.gemm_group([gemm.make_desc(a, b, M, Nn, K, N.EPI_FWD_XENT, bias=bias, out=dl, outT=dlT, outf=logits,
                                    labels=labels, metrics_ptr=metrics.data_ptr())])
    torch.cuda.synchronize()
    z = (a[:, :K].float() @ b[:, :K].float().t() + bias).requires_grad_(True)
    loss = torch.nn.functional.cross_entropy(z, labels)
    loss.backward()
    acc = (z.argmax(1) == labels).float().mean().item()
    assert abs(metrics[0].item() - loss.item()) < 1e-4
    assert abs(metrics[1].item() - acc) < 1e-6
    assert torch.allclose(logits[:, :Nn], z.detach(), rtol=1e-4, atol=1e-4)
    assert torch.allclose(dl[:, :Nn].float(), z.grad, atol=2e-5 + 4e-3 * z.grad.abs().max().item())
    assert torch.equal(dlT[:, :M].t().contiguous(), dl[:, :Nn].contiguous())
    # stand-alone kernel on the same logits
    m2 = torch.zeros(2, device="cuda")
    dl2 = torch.zeros_like(dl)
    gemm.softmax_xent(logits, labels, Nn, dl2, None, m2)
    torch.cuda.synchronize()
    assert abs(m2[0].item() - loss.item()) < 1e-4 and abs(m2[1].item() - acc) < 1e-6
    # different summation order (warp-parallel vs one thread per row): equal up to one bf16 ulp
    assert torch.allclose(dl2.float(), dl.float(), rtol=1e-2, atol=1e-6)


def test_grouped_launch_dgrad_plus_wgrad(ops):
    N, gemm = ops
    B, fin, fout = 128, 200, 100
    dz = _mk(B, fout, ld=104, seed=11)
    dzT = torch.zeros(fout, B, device="cuda", dtype=torch.bfloat16)
    dzT[:, :] = dz[:, :fout].t()
    w = _mk(fin, fout, ld=104, seed=12)
    aT = torch.ones(fin + 1, B, device="cuda", dtype=torch.bfloat16)
    aT[:fin] = _mk(fin, B, seed=13)
    dprev = torch.zeros(B, 200, device="cuda", dtype=torch.bfloat16)
    dw = torch.zeros(fin + 1, fout, device="cuda")
    gemm.gemm_group([gemm.make_desc(dz, w, B, fin, fout, N.EPI_DGRAD, out=dprev),
                     gemm.make_desc(aT, dzT, fin + 1, fout, B, N.EPI_WGRAD, outf=dw)])
    torch.cuda.synchronize()
    r1 = dz[:, :fout].float() @ w[:, :fout].float().t()
    r2 = aT.float() @ dzT.float().t()
    assert torch.allclose(dprev.float(), r1, rtol=1e-2, atol=1e-2 * r1.abs().max().item())
    assert torch.allclose(dw, r2, rtol=1e-4, atol=1e-3)
    assert torch.allclose(dw[fin], dz[:, :fout].float().sum(0), rtol=1e-4, atol=1e-3)  # ones row -> bias gradient


def test_prep_input_and_device_pool(ops):
    N, gemm = ops
    P, B, F = 5, 128, 784
    xs = torch.rand(P, B, F, device="cuda")
    ys = torch.randint(0, 10, (P, B), device="cuda")
    xb = torch.zeros(B, F, device="cuda", dtype=torch.bfloat16)
    xT = torch.zeros(F + 1, B, device="cuda", dtype=torch.bfloat16)
    lab = torch.zeros(B, device="cuda", dtype=torch.int64)
    step = torch.tensor([13], device="cuda", dtype=torch.int64)
    gemm.prep_input(xs, xb, xT, labels_pool=ys, labels_out=lab, step_ptr=step.data_ptr())
    torch.cuda.synchronize()
    j = 13 % P
    assert torch.equal(xb, xs[j].to(torch.bfloat16))
    assert torch.equal(xT[:F], xs[j].to(torch.bfloat16).t())
    assert torch.equal(lab, ys[j])
    gemm.prep_input(xs[0], xb, None)
    torch.cuda.synchronize()
    assert torch.equal(xb, xs[0].to(torch.bfloat16))


def _store(shapes, hyper=None, dev="cuda"):
    from tensorflow_train_distributed_b200.parallel.param_store import GpuParamStore, Hyper
    from tensorflow_train_distributed_b200.parallel.sharding import LayerGroup, ParamLayout
    groups = [LayerGroup(f"l{i}", r, c, i < len(shapes) - 1, (r, c)) for i, (r, c) in enumerate(shapes)]
    return GpuParamStore(ParamLayout(groups), hyper or Hyper(), torch.device(dev), need_w_rm_first=True)


def test_init_is_truncated_he_normal(ops):
    s = _store([(784, 200), (200, 100), (100, 50), (50, 25), (25, 10)])
    s.init(42)
    torch.cuda.synchronize()
    k1 = s.w[:784 * 200]
    sigma = (2.6 / 784) ** 0.5
    assert abs(k1.std().item() - 0.8796 * sigma) / (0.8796 * sigma) < 0.03  # std of N(0,1) truncated at 2
    assert abs(k1.mean().item()) < 3 * sigma / (784 * 200) ** 0.5 * 2
    assert k1.abs().max().item() <= 2.0001 * sigma
    assert s.bias(0).abs().max().item() == 0
    assert torch.equal(s.w, s.ema)
    assert torch.equal(s.w_t[0][:, :784].float(), k1.view(784, 200).t().to(torch.bfloat16).float())
    assert torch.equal(s.w_rm[1][:, :100].float(), s.w[s.layout.groups[1].offset:][:200 * 100].view(200, 100).to(torch.bfloat16).float())
    s2 = _store([(784, 200), (200, 100), (100, 50), (50, 25), (25, 10)])
    s2.init(42)
    assert torch.equal(s.w, s2.w)  # deterministic in (seed, index): every replica computes the same init
    s2.init(43)
    assert not torch.equal(s.w, s2.w)


@pytest.mark.parametrize("reg", ["none", "l1", "l2"])
def test_update_matches_closed_form(ops, reg):
    from tensorflow_train_distributed_b200.parallel.param_store import Hyper
    from tensorflow_train_distributed_b200.utils.schedule import ema_decay, exponential_decay
    h = Hyper(lr0=0.01, decay_rate=0.96, decay_steps=468, regularizer=reg, reg_scale=0.01, ema_decay=0.99)
    s = _store([(784, 200), (200, 100), (25, 10)], h)
    s.init(7)
    s.set_global_step(1000)
    g = torch.randn_like(s.w) * 0.1
    s.grads.copy_(g)
    w0, e0 = s.w.clone(), s.ema.clone()
    s.apply_update()
    torch.cuda.synchronize()
    lr = exponential_decay(0.01, 1000, 468, 0.96, True)
    regmask = torch.zeros_like(w0)
    for gr in s.layout.groups:
        if gr.regularize:
            regmask[gr.offset:gr.offset + gr.rows * gr.cols] = 1
    valid = torch.zeros_like(w0, dtype=torch.bool)
    for gr in s.layout.groups:
        valid[gr.offset:gr.offset + gr.numel] = True
    term = {"none": 0 * w0, "l1": 0.01 * torch.sign(w0), "l2": 0.01 * w0}[reg] * regmask
    wref = w0 - lr * (g + term)
    d = ema_decay(0.99, 1001)
    eref = e0 - (1 - d) * (e0 - wref)
    assert torch.allclose(s.w[valid], wref[valid], rtol=1e-6, atol=1e-7)
    assert torch.allclose(s.ema[valid], eref[valid], rtol=1e-6, atol=1e-7)
    st = s.read_state()
    assert st.global_step == 1001 and st.local_step == 1
    assert abs(st.lr_last - lr) < 1e-9
    g0 = s.layout.groups[0]
    assert torch.equal(s.w_t[0][:, :784].float(), s.w[:g0.rows * g0.cols].view(784, 200).t().to(torch.bfloat16).float())
    # metric accumulators are published and cleared by the same launch
    s.state.view(torch.float32)[4:6] = torch.tensor([1.5, 0.25], device="cuda")
    s.apply_update()
    torch.cuda.synchronize()
    st = s.read_state()
    assert (st.metrics_out[0], st.metrics_out[1]) == (1.5, 0.25) and st.metrics_acc[0] == 0.0


def test_update_respects_owner_range(ops):
    s = _store([(784, 200), (25, 10)])
    s.init(1)
    s.grads.fill_(1.0)
    w0 = s.w.clone()
    s.lo, s.hi = 1000, 5000
    s.apply_update()
    torch.cuda.synchronize()
    changed = (s.w != w0)
    assert changed[1000:5000].all() and not changed[:1000].any() and not changed[5000:].any()


def test_pull_fused_b_operand_reads_each_k_range_from_its_owner(ops):
    """K3 on one GPU: every "replica" buffer holds valid data only in the K range it owns (NaN elsewhere);
    the GEMM must stitch the B operand together from the right owners."""
    N, gemm = ops
    M, Nn, K = 128, 200, 784
    a, b = _mk(M, K, scale=0.05, seed=21), _mk(Nn, K, seed=22)
    bounds = [0, 128, 256, 320, 448, 576, 704, 784, 784]
    peers = []
    for o in range(8):
        pb = torch.full_like(b, float("nan"))
        pb[:, bounds[o]:bounds[o + 1]] = b[:, bounds[o]:bounds[o + 1]]
        peers.append(pb)
    poisoned_local = torch.full_like(b, float("nan"))
    out = torch.zeros(M, 200, device="cuda")
    d = gemm.make_desc(a, poisoned_local, M, Nn, K, N.EPI_FWD, outf=out, b_peers=[p.data_ptr() for p in peers],
                       k_bounds=bounds)
    gemm.gemm_group([d])
    torch.cuda.synchronize()
    ref = a[:, :K].float() @ b[:, :K].float().t()
    assert torch.isfinite(out).all()
    assert torch.allclose(out, ref, rtol=1e-4, atol=1e-3 * K ** 0.5)


@pytest.mark.parametrize("mode", ["fwd", "dgrad"])
def test_weight_stationary_orientation_matches_plain_orientation(ops, mode):
    """Wide layers run with exchanged operands (tile rows = features); results and dropout masks must not care."""
    N, gemm = ops
    M, Nn, K, rate = 128, 2048, 512, 0.1
    a, b = _mk(M, K, scale=0.05, seed=31), _mk(Nn, K, seed=32)
    bias = torch.randn(Nn, device="cuda")
    aux = _mk(M, Nn, seed=33)
    step = torch.tensor([3], device="cuda", dtype=torch.int64)
    res = []
    for no_swap in (True, False):
        out = torch.zeros(M, Nn, device="cuda", dtype=torch.bfloat16)
        outT = torch.zeros(Nn + 1, M, device="cuda", dtype=torch.bfloat16)
        outf = torch.zeros(M, Nn, device="cuda") if mode == "fwd" else None
        if mode == "fwd":
            d = gemm.make_desc(a, b, M, Nn, K, N.EPI_FWD, act=N.ACT_ELU, drop_rate=rate, layer_id=2, bias=bias, out=out,
                               outT=outT, outf=outf)
        else:
            d = gemm.make_desc(a, b, M, Nn, K, N.EPI_DGRAD, act=N.ACT_ELU, drop_rate=rate, layer_id=2, aux=aux, out=out,
                               outT=outT)
        gemm.gemm_group([d], seed=99, step_ptr=step.data_ptr(), no_swap=no_swap, no_cluster=True)
        torch.cuda.synchronize()
        res.append((out.clone(), outT.clone(), None if outf is None else outf.clone()))
    (o0, t0, f0), (o1, t1, f1) = res
    assert torch.equal(o0 != 0, o1 != 0)                       # identical dropout pattern
    assert torch.allclose(o0.float(), o1.float(), rtol=1e-2, atol=1e-3)
    assert torch.equal(t1[:Nn].t().contiguous(), o1)           # both copies agree in the exchanged orientation too
    if mode == "fwd":
        ref = torch.nn.functional.elu(a.float() @ b.float().t() + bias)
        assert torch.allclose(f1, ref, rtol=1e-3, atol=1e-3) and torch.allclose(f0, f1, rtol=1e-5, atol=1e-5)
        assert abs((o1 == 0).float().mean().item() - rate) < 0.02
