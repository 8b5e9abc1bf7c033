"""Python face of the grouped tcgen05 GEMM (``csrc/gemm_tcgen05.cu``).

All matrices are bf16 with the reduction dimension contiguous ("K-major"), so a layer needs

* forward   ``Z[B,out] = A[B,in] · Wt[out,in]^T``        (Wt = transposed compute copy)
* dgrad     ``dA[B,in] = dZ[B,out] · W[in,out]^T``       (W = TF-layout compute copy)
* wgrad     ``dW[in+1,out] = At[in+1,B] · dZt[out,B]^T``  (transposed activations; row ``in`` of At is
  all ones, which makes row ``in`` of dW the bias gradient)

which is why every epilogue can also emit the transposed copy of what it produces.
Reference: forward_propagation / loss / minimize in distribute_training.py:39-63, :66-87, :152.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import _native as N


def _ld(t: torch.Tensor) -> int:
    assert t.dim() == 2 and t.stride(1) == 1, "matrix must be 2-D with unit inner stride"
    return t.stride(0)


def make_desc(a: torch.Tensor, b: torch.Tensor, M: int, Nn: int, K: int, mode: int, *, act: int = N.ACT_NONE,
              drop_rate: float = 0.0, layer_id: int = 0, bias: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None, outT: Optional[torch.Tensor] = None,
              outf: Optional[torch.Tensor] = None, aux: Optional[torch.Tensor] = None, scale: float = 1.0,
              labels: Optional[torch.Tensor] = None, metrics_ptr: Optional[int] = None,
              block_n: int = 0, b_peers: Optional[Sequence[int]] = None,
              k_bounds: Optional[Sequence[int]] = None, upd_w: Optional[torch.Tensor] = None,
              upd_ema: Optional[torch.Tensor] = None, upd_bias: Optional[torch.Tensor] = None,
              upd_regularize: bool = False, a_slot_rows: int = 0, a_slots: int = 0) -> N.GemmDesc:
    """Describe one GEMM problem. ``a`` is [>=M, lda] and ``b`` is [>=N, ldb] bf16 storage; M/N/K are the
    logical extents (TMA zero-fills everything outside them)."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    d = N.GemmDesc()
    d.a, d.lda = a.data_ptr(), _ld(a)
    d.b, d.ldb = b.data_ptr(), _ld(b)
    d.M, d.N, d.K = M, Nn, K
    d.block_n = block_n
    d.mode, d.act = mode, act
    d.drop_rate, d.layer_id = float(drop_rate), layer_id
    d.bias = N.ptr(bias)
    if bias is not None:
        assert bias.dtype == torch.float32
    if out is not None:
        assert out.dtype == torch.bfloat16
        d.out, d.ldo = out.data_ptr(), _ld(out)
    if outT is not None:
        assert outT.dtype == torch.bfloat16
        d.outT, d.ldt = outT.data_ptr(), _ld(outT)
    if outf is not None:
        assert outf.dtype == torch.float32
        d.outf, d.ldf = outf.data_ptr(), _ld(outf)
    if aux is not None:
        assert aux.dtype == torch.bfloat16
        d.aux, d.ldaux = aux.data_ptr(), _ld(aux)
    d.scale = float(scale)
    if labels is not None:
        assert labels.dtype == torch.int64
        d.labels = labels.data_ptr()
    d.metrics = metrics_ptr
    if upd_w is not None:  # EPI_WGRAD_UPDATE: fp32 master / shadow [M, N] views of the flat buffers, fp32 bias copy
        assert upd_w.dtype == torch.float32 and upd_w.stride(1) == 1
        d.upd_w, d.ldf = upd_w.data_ptr(), upd_w.stride(0)
        if outf is not None:
            assert outf.stride(0) == upd_w.stride(0)
        if upd_ema is not None:
            assert upd_ema.dtype == torch.float32 and upd_ema.stride(0) == upd_w.stride(0)
            d.upd_ema = upd_ema.data_ptr()
        d.upd_bias = N.ptr(upd_bias)
        d.upd_regularize = 1 if upd_regularize else 0
    d.a_slot_rows, d.a_slots = int(a_slot_rows), int(a_slots)
    if b_peers:
        # pull-fused B: k range o of the operand is read from replica o's peer-mapped copy (K3)
        assert k_bounds is not None and len(k_bounds) == len(b_peers) + 1 and len(b_peers) <= N.MAX_PEERS
        for o, ptr in enumerate(b_peers):
            d.b_peer[o] = ptr
        for o, kb in enumerate(k_bounds):
            d.k_bound[o] = int(kb)
        d.n_b_peers = len(b_peers)
    return d


def gemm_group(descs: Sequence[N.GemmDesc], *, seed: int = 0, step_ptr: Optional[int] = None, pdl: bool = False,
               stream: Optional[int] = None, no_cluster: bool = False, no_swap: bool = False) -> None:
    """Launch 1..4 problems as one kernel on the current stream. ``no_cluster`` / ``no_swap`` switch off the
    A-multicast clusters and the weight-stationary operand exchange (tests compare both code paths)."""
    lib = N.load()
    arr = (N.GemmDesc * len(descs))(*descs)
    flags = (1 if pdl else 0) | (2 if no_cluster else 0) | (4 if no_swap else 0)
    rc = lib.ttd_gemm_group(arr, len(descs), C.c_uint64(seed & (2**64 - 1)), step_ptr, flags,
                            stream if stream is not None else N.cur_stream())
    N.check(rc, "ttd_gemm_group")


def softmax_xent(logits: torch.Tensor, labels: torch.Tensor, C_: int, dlogits: Optional[torch.Tensor],
                 dlogitsT: Optional[torch.Tensor], metrics: torch.Tensor, scale: float = 1.0) -> None:
    lib = N.load()
    B = labels.numel()
    rc = lib.ttd_softmax_xent(logits.data_ptr(), _ld(logits), labels.data_ptr(), B, C_,
                              N.ptr(dlogits), _ld(dlogits) if dlogits is not None else 0,
                              N.ptr(dlogitsT), _ld(dlogitsT) if dlogitsT is not None else 0,
                              float(scale), metrics.data_ptr(), N.cur_stream())
    N.check(rc, "ttd_softmax_xent")


def prep_input(x: torch.Tensor, xb: Optional[torch.Tensor], xT: Optional[torch.Tensor], *,
               labels_pool: Optional[torch.Tensor] = None, labels_out: Optional[torch.Tensor] = None,
               step_ptr: Optional[int] = None, xf: Optional[torch.Tensor] = None, pdl: bool = False) -> None:
    """Stage one fp32 batch as bf16 (+ transposed). ``x`` is [B, F], or a device-resident pool [P, B, F] from
    which batch ``*step_ptr % P`` is taken (its labels go to ``labels_out``)."""
    lib = N.load()
    assert x.dtype == torch.float32 and x.is_contiguous()
    pool = 1
    if x.dim() == 3:
        pool, B, F = x.shape
    else:
        B, F = x.shape
    rc = lib.ttd_prep_input_ex(x.data_ptr(), B, F, N.ptr(xb), _ld(xb) if xb is not None else 0,
                               N.ptr(xT), _ld(xT) if xT is not None else 0, N.ptr(labels_pool), N.ptr(labels_out),
                               pool, step_ptr, N.ptr(xf), 1 if pdl else 0, N.cur_stream())
    N.check(rc, "ttd_prep_input")


class GemmChain:
    """A fixed list of small dependent GEMM problems executed by ONE CTA per launch (``gemm_chain_kernel``)."""

    def __init__(self, descs: Sequence[N.GemmDesc], device: torch.device):
        lib = N.load()
        self.n = len(descs)
        nbytes = int(lib.ttd_gemm_chain_problem_bytes()) * self.n
        self.buf = torch.zeros(nbytes + 64, dtype=torch.uint8, device=device)
        self.ptr = (self.buf.data_ptr() + 63) & ~63
        arr = (N.GemmDesc * self.n)(*descs)
        N.check(lib.ttd_gemm_chain_build(arr, self.n, self.ptr, N.cur_stream()), "ttd_gemm_chain_build")

    def launch(self, *, seed: int = 0, step_ptr: Optional[int] = None, pdl: bool = False) -> None:
        lib = N.load()
        N.check(lib.ttd_gemm_chain_launch(self.ptr, self.n, C.c_uint64(seed & (2**64 - 1)), step_ptr,
                                          1 if pdl else 0, N.cur_stream()), "ttd_gemm_chain_launch")


class GemmNet:
    """All GEMMs of a training step as ONE launch of co-resident CTAs (``gemm_net_kernel``): every tile is a task,
    dependent problems hand over through arrival counters in global memory instead of kernel boundaries.

    ``deps[i]`` lists the (earlier) problems whose output problem ``i`` reads."""

    MAX_TASKS = 8

    def __init__(self, descs: Sequence[N.GemmDesc], deps: Sequence[Sequence[int]], device: torch.device):
        lib = N.load()
        self.n = len(descs)
        self.grid = torch.cuda.get_device_properties(device).multi_processor_count
        nbytes = int(lib.ttd_gemm_net_bytes(self.n, self.grid, self.MAX_TASKS))
        self.buf = torch.zeros(nbytes + 64, dtype=torch.uint8, device=device)
        self.ptr = (self.buf.data_ptr() + 63) & ~63
        arr = (N.GemmDesc * self.n)(*descs)
        masks = (C.c_uint32 * self.n)(*[sum(1 << d for d in dd) for dd in deps])
        rc = lib.ttd_gemm_net_build(arr, self.n, masks, self.grid, self.MAX_TASKS, self.ptr, N.cur_stream())
        if rc <= 0:
            N.check(rc if rc < 0 else -1, "ttd_gemm_net_build")
        self.ctas = rc

    def launch(self, *, seed: int = 0, step_ptr: Optional[int] = None, pdl: bool = False,
               hyper: Optional[N.Hyper] = None, lr_slot: Optional[int] = None, state_ptr: Optional[int] = None,
               finalize: int = 0) -> None:
        """``hyper`` / ``lr_slot`` / ``state_ptr`` serve EPI_WGRAD_UPDATE problems (the optimiser step in the
        weight-gradient epilogue); ``finalize`` bit0: the last CTA ends the training step, bit1: ++global_step."""
        lib = N.load()
        N.check(lib.ttd_gemm_net_launch_ex(self.ptr, self.n, self.grid, self.MAX_TASKS, self.ctas,
                                           C.c_uint64(seed & (2**64 - 1)), step_ptr, 1 if pdl else 0,
                                           C.byref(hyper) if hyper is not None else None, lr_slot, state_ptr,
                                           int(finalize), N.cur_stream()),
                "ttd_gemm_net_launch")


class MlpColumn:
    """Forward + loss + data gradients of a narrow MLP as ONE launch (``mlp_column_kernel``): a CTA owns 16 examples
    through every layer, activations stay in shared memory. ``desc`` is a filled :class:`N.ColumnDesc`."""

    def __init__(self, desc: N.ColumnDesc):
        self.lib = N.load()
        self._desc = desc  # keeps the pointers' owners findable while debugging; the handle has its own copy
        self.handle = self.lib.ttd_mlp_column_create(C.byref(desc))
        if not self.handle:
            raise N.NativeError("ttd_mlp_column_create failed: " + self.lib.ttd_last_error().decode(errors="replace"))

    def launch(self, *, seed: int = 0, pdl: bool = False) -> None:
        N.check(self.lib.ttd_mlp_column_launch(self.handle, C.c_uint64(seed & (2**64 - 1)), 1 if pdl else 0,
                                               N.cur_stream()), "ttd_mlp_column_launch")

    def close(self) -> None:
        if self.handle:
            self.lib.ttd_mlp_column_destroy(self.handle)
            self.handle = None

    def __del__(self):  # noqa: D105
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def column_fits(input_dim: int, layer_sizes: Sequence[int]) -> bool:
    lib = N.load()
    arr = (C.c_int32 * len(layer_sizes))(*[int(x) for x in layer_sizes])
    return bool(lib.ttd_mlp_column_fits(int(input_dim), arr, len(layer_sizes)))
