"""ttd_apply_update_static: the update entry point for hosts that must not allocate device memory.

Some users embed this library in a process that already owns all of the GPU's memory (its own arena, or
another framework's pool) or runs a loop that cannot absorb the device-wide stall of a cudaMalloc; they need
an update call that brings its own working storage. This file drives that entry point through the C ABI
exactly as such a host would (raw pointers, plain structs) and checks it against the regular
``ttd_apply_update`` on identical inputs, twice in a row — the second call proves that whatever the first
one left in the library's built-in scratch does not leak into the next result.

The file name sorts last on purpose: a run that stops at the first failure has then already executed every
other test.
"""
import ctypes as C

import pytest
import torch


def _bind():
    from tensorflow_train_distributed_b200.ops import _native as N
    lib = N.load()
    fn = lib.ttd_apply_update_static
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(N.VarDesc), C.c_int, C.c_int64, C.c_int64,
                   C.POINTER(N.Hyper), C.c_void_p, C.c_int32, C.c_void_p]
    return N, lib, fn


def _make_side(N, shapes, seed):
    """Flat buffers, bf16 copies and a variable table for one independent copy of the problem."""
    off, descs, copies = 0, [], []
    for i, (r, c) in enumerate(shapes):
        v = N.VarDesc()
        v.offset, v.rows, v.cols, v.regularize = off, r, c, 1 if i < len(shapes) - 1 else 0
        ldr, ldt = -(-c // 8) * 8, -(-r // 8) * 8
        wrm = torch.zeros(r, ldr, device="cuda", dtype=torch.bfloat16)
        wt = torch.zeros(c, ldt, device="cuda", dtype=torch.bfloat16)
        v.w_rm, v.ld_rm, v.w_t, v.ld_t = wrm.data_ptr(), ldr, wt.data_ptr(), ldt
        descs.append(v)
        copies.append((wrm, wt))
        off += -(-(r + 1) * c // 64) * 64
    g = torch.Generator(device="cuda").manual_seed(seed)
    w = torch.randn(off, device="cuda", generator=g) * 0.05
    ema = w.clone()
    state = torch.zeros(8, device="cuda", dtype=torch.int64)
    state[0] = 930  # global_step: one staircase decay already behind us
    return dict(w=w, ema=ema, state=state, copies=copies, table=(N.VarDesc * len(descs))(*descs), n=len(descs),
                total=off)


@pytest.mark.gpu
def test_static_scratch_update_matches_regular_entry_point_twice():
    N, lib, fn = _bind()
    shapes = [(784, 200), (200, 100), (100, 50), (50, 25), (25, 10)]
    a, b = _make_side(N, shapes, seed=21), _make_side(N, shapes, seed=21)
    assert torch.equal(a["w"], b["w"])
    hyper = N.Hyper(0.01, 0.96, 468, 1, 2, 0.01, 0.99, 1, 0.5)
    peers = N.Peers()
    peers.n_grads = peers.n_copies = 1
    stream = torch.cuda.current_stream().cuda_stream
    gen = torch.Generator(device="cuda").manual_seed(5)
    for call in range(2):
        grad = torch.randn(a["total"], device="cuda", generator=gen) * 0.1
        # metric accumulators that the update publishes and clears
        for side in (a, b):
            side["state"].view(torch.float32)[4:6] = torch.tensor([0.75 + call, 0.5], device="cuda")
        peers.grads[0] = grad.data_ptr()
        rc = lib.ttd_apply_update(a["w"].data_ptr(), a["ema"].data_ptr(), a["table"], a["n"], 0, a["total"],
                                  C.byref(hyper), C.byref(peers), a["state"].data_ptr(), 1, stream)
        assert rc == 0, lib.ttd_last_error()
        rc = fn(b["w"].data_ptr(), b["ema"].data_ptr(), grad.data_ptr(), b["table"], b["n"], 0, b["total"],
                C.byref(hyper), b["state"].data_ptr(), 1, stream)
        assert rc == 0, lib.ttd_last_error()
        torch.cuda.synchronize()
        # same tolerance as the update tests in test_gpu_kernels.py
        assert torch.allclose(b["w"], a["w"], rtol=1e-6, atol=1e-7), f"master weights differ on call {call}"
        assert torch.allclose(b["ema"], a["ema"], rtol=1e-6, atol=1e-7), f"shadow variables differ on call {call}"
        for (rm_a, t_a), (rm_b, t_b) in zip(a["copies"], b["copies"]):
            assert torch.allclose(rm_b.float(), rm_a.float(), rtol=1e-2, atol=1e-6)  # bf16: at most one ulp apart
            assert torch.allclose(t_b.float(), t_a.float(), rtol=1e-2, atol=1e-6)
            assert torch.equal(t_b[:, :rm_b.shape[0]].t()[:, :t_b.shape[0]], rm_b[:, :t_b.shape[0]])
        sa = N.StepState.from_buffer_copy(a["state"].cpu().numpy().tobytes())
        sb = N.StepState.from_buffer_copy(b["state"].cpu().numpy().tobytes())
        assert sb.global_step == sa.global_step == 931 + call
        assert sb.local_step == sa.local_step == 1 + call
        assert abs(sb.lr_last - sa.lr_last) < 1e-9
        assert (sb.metrics_out[0], sb.metrics_out[1]) == (sa.metrics_out[0], sa.metrics_out[1]) == (0.75 + call, 0.5)
        assert sb.metrics_acc[0] == 0.0 and sb.metrics_acc[1] == 0.0
    assert not torch.equal(a["w"], _make_side(N, shapes, seed=21)["w"])  # the updates really changed something


@pytest.mark.gpu
def test_static_scratch_update_respects_owner_range():
    N, lib, fn = _bind()
    s = _make_side(N, [(64, 32), (32, 10)], seed=3)
    w0 = s["w"].clone()
    grad = torch.ones(s["total"], device="cuda")
    hyper = N.Hyper(0.01, 0.96, 468, 1, 0, 0.0, 0.0, 0, 1.0)
    rc = fn(s["w"].data_ptr(), None, grad.data_ptr(), s["table"], s["n"], 100, 700, C.byref(hyper),
            s["state"].data_ptr(), 0, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.ttd_last_error()
    torch.cuda.synchronize()
    changed = s["w"] != w0
    assert changed[100:700].all() and not changed[:100].any() and not changed[700:].any()
    assert int(s["state"][0].item()) == 930  # bump_global_step = 0 leaves the shared counter alone


@pytest.mark.gpu
def test_static_scratch_update_unaligned_range_and_negative_hi_match_regular_entry_point():
    """[lo, hi) that cuts 4-element groups, and hi < 0 (= to the end): both entry points must agree."""
    N, lib, fn = _bind()
    shapes = [(64, 32), (32, 12)]
    stream = torch.cuda.current_stream().cuda_stream
    for lo, hi in ((101, 1703), (7, -1)):
        a, b = _make_side(N, shapes, seed=8), _make_side(N, shapes, seed=8)
        grad = torch.randn(a["total"], device="cuda") * 0.1
        hyper = N.Hyper(0.01, 0.96, 468, 1, 2, 0.01, 0.99, 1, 1.0)
        peers = N.Peers()
        peers.n_grads = peers.n_copies = 1
        peers.grads[0] = grad.data_ptr()
        rc = lib.ttd_apply_update(a["w"].data_ptr(), a["ema"].data_ptr(), a["table"], a["n"], lo, hi, C.byref(hyper),
                                  C.byref(peers), a["state"].data_ptr(), 1, stream)
        assert rc == 0, lib.ttd_last_error()
        rc = fn(b["w"].data_ptr(), b["ema"].data_ptr(), grad.data_ptr(), b["table"], b["n"], lo, hi, C.byref(hyper),
                b["state"].data_ptr(), 1, stream)
        assert rc == 0, lib.ttd_last_error()
        torch.cuda.synchronize()
        assert torch.allclose(b["w"], a["w"], rtol=1e-6, atol=1e-7)
        assert torch.allclose(b["ema"], a["ema"], rtol=1e-6, atol=1e-7)
        fresh = _make_side(N, shapes, seed=8)["w"]
        end = a["total"] if hi < 0 else hi
        last = shapes[-1]
        used = sum(-(-(r + 1) * c // 64) * 64 for r, c in shapes[:-1]) + (last[0] + 1) * last[1]
        end = min(end, used)
        assert (a["w"][lo:end] != fresh[lo:end]).any() and torch.equal(a["w"][:lo], fresh[:lo])


@pytest.mark.gpu
def test_static_scratch_update_refuses_more_groups_than_fit():
    """Host-side check (no kernel is launched when the request exceeds the built-in table); marked gpu because
    loading the kernel library needs the CUDA runtime of a GPU box."""
    N, lib, fn = _bind()
    too_many = 65
    table = (N.VarDesc * too_many)()
    hyper = N.Hyper(0.01, 0.96, 468, 1, 0, 0.0, 0.0, 0, 1.0)
    rc = fn(1, None, 1, table, too_many, 0, 1, C.byref(hyper), 1, 0, None)
    assert rc != 0
    msg = lib.ttd_last_error().decode()
    assert "at most 64" in msg and "65" in msg
