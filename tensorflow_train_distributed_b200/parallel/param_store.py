"""GPU parameter store: the "ps" of the reference as device memory.

What lives here replaces the ps job of the reference (``server.join()``, distribute_training.py:179-181, and
the ``ApplyGradientDescent`` / ``AssignAdd(global_step)`` ops TF places on it, :150-152):

* ``w``      flat fp32 master parameters, one ``[rows+1, cols]`` group per layer (kernel + bias); a replica's
             copy is authoritative only inside the range it owns
* ``ema``    flat fp32 shadow variables (``<var>/ExponentialMovingAverage``)
* ``grads``  flat fp32 gradient buffer — the wgrad GEMMs write straight into it
* ``copies`` compute copies refreshed by the update kernel on *every* replica (that is the parameter pull):
             bf16 kernels in the two K-major layouts the GEMMs need (``w_t`` forward, ``w_rm`` dgrad) and the
             fp32 biases the forward epilogues read
* ``state``  64-byte device block: global_step, local_step, metric accumulators (never read by the host on the
             step path)

With several GPUs, ``grads``, ``w``, the copies and a signal pad live in ONE symmetric (peer-mapped)
allocation, so the address of anything on replica *r* is the local address plus one constant.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Optional

import numpy as np
import torch

from ..ops import _native as N
from .sharding import EMA_SUFFIX, GLOBAL_STEP, FlatPlan, ParamLayout, flat

FLAG_SLOTS = 148 * 4  # one signal row per CTA of a barrier-mode launch (<= 4 co-resident CTAs per SM)


def _pad8(n: int) -> int:
    return -(-n // 8) * 8


def _align(n: int, a: int) -> int:
    return -(-n // a) * a


@dataclass
class Hyper:
    """Optimiser hyper-parameters evaluated on the device (reference :12-14, :39, :120-122, :136-140)."""
    lr0: float = 0.01
    decay_rate: float = 0.96
    decay_steps: int = 468
    staircase: bool = True
    regularizer: str = "l2"
    reg_scale: float = 0.01
    ema_decay: float = 0.99  # <= 0 disables
    ema_warmup: bool = True
    grad_scale: float = 1.0

    def to_native(self) -> N.Hyper:
        return N.Hyper(self.lr0, self.decay_rate, int(self.decay_steps), 1 if self.staircase else 0,
                       N.REG_CODES[self.regularizer], self.reg_scale, self.ema_decay,
                       1 if self.ema_warmup else 0, self.grad_scale)


class GpuParamStore:
    def __init__(self, layout: ParamLayout, hyper: Hyper, device: torch.device, *, comm=None,
                 need_w_rm_first: bool = False, pull_fused_first: bool = False, tail_groups=()):
        """``comm`` is a :class:`~.symmetric.SymmetricComm` for multi-GPU runs (None: single GPU)."""
        self.layout, self.hyper, self.device, self.comm = layout, hyper, device, comm
        # K3: the first layer's transposed weights are not broadcast after an update; every replica's first
        # GEMM fetches each k-block from its owner by TMA over NVLink instead
        self.pull_fused_first = bool(pull_fused_first and comm is not None)
        import os as _os
        # NVLS (in-switch reduction / replication) pays off once several peers are involved; with two GPUs plain
        # peer loads and stores are ~1.6x faster (measured: 533 vs 850 us for the 73.6 M-parameter push)
        nvls_env = _os.environ.get("TTD_NVLS", "auto")
        self.use_multicast = (nvls_env == "1") or (nvls_env == "auto" and comm is not None and comm.world_size > 2)
        # groups served by the fused MLP tail kernel: fp32 compute copies (both layouts) instead of bf16 ones
        self.tail_groups = frozenset(tail_groups)
        self.lib = N.load()
        self.world = comm.world_size if comm is not None else 1
        self.rank = comm.rank if comm is not None else 0
        total = layout.total

        # ---- carve one arena: grads | w | bf16 copies | fp32 bias copies | flags --------------------
        sizes = []
        for i, g in enumerate(layout.groups):
            tail = i in self.tail_groups
            want_rm = (need_w_rm_first or i > 0) and not tail  # first-layer dgrad is never needed (input has no grad)
            sizes.append((g.rows * _pad8(g.cols) if want_rm else 0, 0 if tail else g.cols * _pad8(g.rows), g.cols))
        bf16_elems = sum(_align(a, 128) + _align(b, 128) for a, b, _ in sizes)
        bias_elems = sum(_align(c, 64) for _, _, c in sizes)
        f32_sizes = [(_align(g.rows, 4) * _align(g.cols, 4), _align(g.cols, 4) * _align(g.rows, 4))
                     if i in self.tail_groups else (0, 0) for i, g in enumerate(layout.groups)]
        f32_elems = sum(_align(a, 64) + _align(b, 64) for a, b in f32_sizes)
        o_grads = 0
        o_w = o_grads + total * 4
        o_ema = o_w + total * 4
        o_state = o_ema + total * 4
        o_bf16 = o_state + 256
        o_bias = o_bf16 + _align(bf16_elems * 2, 256)
        o_f32 = o_bias + _align(bias_elems * 4, 256)
        o_flags = o_f32 + _align(f32_elems * 4, 256)
        nbytes = o_flags + FLAG_SLOTS * N.MAX_PEERS * 4
        if comm is not None:
            self._arena = comm.alloc(nbytes)
            raw = self._arena.local
        else:
            self._arena = None
            raw = torch.empty(_align(nbytes, 4096), device=device, dtype=torch.uint8)
        raw.zero_()
        self._raw = raw
        self._flags_off, self._flags_bytes = o_flags, FLAG_SLOTS * N.MAX_PEERS * 4
        self.grads = raw[o_grads:o_grads + total * 4].view(torch.float32)
        self.w = raw[o_w:o_w + total * 4].view(torch.float32)
        copies = raw[o_bf16:o_bf16 + bf16_elems * 2].view(torch.bfloat16)
        biases = raw[o_bias:o_bias + bias_elems * 4].view(torch.float32)
        f32c = raw[o_f32:o_f32 + f32_elems * 4].view(torch.float32)
        self.ema = raw[o_ema:o_ema + total * 4].view(torch.float32)
        self.state = raw[o_state:o_state + 64].view(torch.int64)  # TtdStepState (64 B)
        # this step's learning rate and EMA decay, written by the column kernel for the fused update epilogue
        self.lr_slot = raw[o_state + 64:o_state + 72].view(torch.float32)
        self._off = dict(grads=o_grads, w=o_w, ema=o_ema, state=o_state)

        self.w_rm: List[Optional[torch.Tensor]] = []
        self.w_t: List[torch.Tensor] = []
        self.bias_c: List[torch.Tensor] = []
        self.w_f32: List[Optional[torch.Tensor]] = []
        self.wt_f32: List[Optional[torch.Tensor]] = []
        foff = 0
        for g, (fa, fb) in zip(layout.groups, f32_sizes):
            if fa:
                self.w_f32.append(f32c[foff:foff + fa].view(_align(g.rows, 4), _align(g.cols, 4)))
                foff += _align(fa, 64)
                self.wt_f32.append(f32c[foff:foff + fb].view(_align(g.cols, 4), _align(g.rows, 4)))
                foff += _align(fb, 64)
            else:
                self.w_f32.append(None)
                self.wt_f32.append(None)
        off = boff = 0
        for g, (rm, t, c) in zip(layout.groups, sizes):
            if rm:
                self.w_rm.append(copies[off:off + rm].view(g.rows, _pad8(g.cols)))
                off += _align(rm, 128)
            else:
                self.w_rm.append(None)
            self.w_t.append(copies[off:off + t].view(g.cols, _pad8(g.rows)) if t else None)
            off += _align(t, 128)
            self.bias_c.append(biases[boff:boff + c])
            boff += _align(c, 64)

        # ---- variable table ----------------------------------------------------------------------
        descs = []
        for g, rm, t, b, wf, wtf in zip(layout.groups, self.w_rm, self.w_t, self.bias_c, self.w_f32, self.wt_f32):
            v = N.VarDesc()
            v.offset, v.rows, v.cols = g.offset, g.rows, g.cols
            v.regularize = 1 if g.regularize else 0
            if rm is not None:
                v.w_rm, v.ld_rm = rm.data_ptr(), rm.stride(0)
            if t is not None:
                v.w_t, v.ld_t = t.data_ptr(), t.stride(0)
            if wf is not None:
                v.w_f32, v.ld_f32 = wf.data_ptr(), wf.stride(0)
                v.wt_f32, v.ld_tf32 = wtf.data_ptr(), wtf.stride(0)
            v.bias_c = b.data_ptr()
            if self.pull_fused_first and not descs:
                v.reserved = 1
            descs.append(v)
        self.vars = (N.VarDesc * len(descs))(*descs)
        self.nvars = len(descs)
        self.plan: FlatPlan = flat(layout, self.world)
        self.lo, self.hi = self.plan.range_of(self.rank)
        self._peers = self._build_peers()
        self._async_peers = self._build_async_peers()
        self._hyper_native = hyper.to_native()
        h1 = Hyper(**{**hyper.__dict__, "grad_scale": 1.0})  # async: every gradient is applied on its own
        self._hyper_async = h1.to_native()

    # ------------------------------------------------------------------------------------------
    def set_host_mirror(self, host_ptr: int, slots: int) -> None:
        """Pinned host ring of 64-byte step states: the kernel that ends a step writes its state there."""
        for h in (self._hyper_native, self._hyper_async):
            h.host_mirror = host_ptr
            h.mirror_slots = slots

    def _build_peers(self) -> N.Peers:
        p = N.Peers()
        if self.comm is None:
            p.grads[0] = self.grads.data_ptr()
            p.n_grads = p.n_copies = 1
            return p
        base = self._arena.peer_ptrs
        for r in range(self.world):
            p.grads[r] = base[r]  # the arena starts with the gradient buffer
            p.copy_delta[r] = base[r] - base[self.rank]
            p.flags[r] = base[r] + self._flags_off
        mc = self._arena.multicast_ptr if self.use_multicast else 0
        if mc:  # NVLS: the switch sums the replicas' gradients and replicates the refreshed copies
            p.grads_mc = mc + self._off["grads"]
            p.mc_delta = mc - base[self.rank]
        p.n_grads = p.n_copies = self.world
        p.rank = self.rank
        p.barrier = 1
        # identical on every replica (their CTAs meet pairwise): sized by the average owner range
        # about one 32x32 tile per CTA: the update is latency-bound (NVLink round trips), so spread it wide
        per_owner = -(-self.layout.total // self.world)
        p.grid = min(FLAG_SLOTS, max(1, -(-per_owner // (32 * 32)) + 2 * self.nvars))
        return p

    def _build_async_peers(self) -> N.AsyncPeers:
        a = N.AsyncPeers()
        base = self._arena.peer_ptrs if self.comm is not None else [self._raw.data_ptr()]
        for r in range(self.world):
            a.w_peer[r] = base[r] + self._off["w"]
        for k in range(self.world + 1):
            a.bounds[k] = self.plan.bounds[k]
        a.state0 = base[0] + self._off["state"]
        a.n_owners, a.rank = self.world, self.rank
        return a

    def first_layer_pull_plan(self):
        """(peer addresses of w_t[0], K boundaries) for the pull-fused first GEMM, or (None, None)."""
        if not self.pull_fused_first:
            return None, None
        g0 = self.layout.groups[0]
        kb = [min(g0.rows, max(0, b - g0.offset) // g0.cols) for b in self.plan.bounds]
        kb[-1] = g0.rows
        # owners whose range lies past the first kernel serve nothing: their k range is empty
        ptrs = [self.peer_ptr(self.w_t[0], r) for r in range(self.world)]
        return ptrs, kb

    def peer_ptr(self, tensor: torch.Tensor, rank: int) -> int:
        """Address of replica ``rank``'s counterpart of a tensor that lives in the symmetric arena."""
        if self.comm is None:
            return tensor.data_ptr()
        return tensor.data_ptr() + (self._arena.peer_ptrs[rank] - self._arena.peer_ptrs[self.rank])

    # ---- views ---------------------------------------------------------------------------------
    def bias(self, i: int) -> torch.Tensor:
        """fp32 bias of layer ``i`` as the forward epilogue reads it (the pulled copy, not the master)."""
        return self.bias_c[i]

    def grad_matrix(self, i: int) -> torch.Tensor:
        g = self.layout.groups[i]
        return self.grads[g.offset:g.offset + g.numel].view(g.rows + 1, g.cols)

    def master_matrix(self, i: int) -> torch.Tensor:
        """fp32 master of group ``i`` as the [rows+1, cols] matrix (kernel rows, then the bias row)."""
        g = self.layout.groups[i]
        return self.w[g.offset:g.offset + g.numel].view(g.rows + 1, g.cols)

    def ema_matrix(self, i: int) -> torch.Tensor:
        g = self.layout.groups[i]
        return self.ema[g.offset:g.offset + g.numel].view(g.rows + 1, g.cols)

    @property
    def hyper_native(self) -> N.Hyper:
        return self._hyper_native

    @property
    def global_step_ptr(self) -> int:
        return self.state.data_ptr() + N.OFF_GLOBAL_STEP

    @property
    def state_ptr(self) -> int:
        return self.state.data_ptr()

    @property
    def local_step_ptr(self) -> int:
        return self.state.data_ptr() + N.OFF_LOCAL_STEP

    @property
    def metrics_acc_ptr(self) -> int:
        return self.state.data_ptr() + N.OFF_METRICS_ACC

    # ---- operations ----------------------------------------------------------------------------
    def init(self, seed: int) -> None:
        """He init of every kernel, zero biases, shadows = values (reference :49; chief-seeded so every replica
        computes identical parameters without a broadcast)."""
        N.check(self.lib.ttd_init_params(self.w.data_ptr(), self.ema.data_ptr(), self.vars, self.nvars,
                                         C.c_uint64(seed), N.cur_stream()), "ttd_init_params")

    def refresh_copies(self) -> None:
        N.check(self.lib.ttd_refresh_copies(self.w.data_ptr(), self.vars, self.nvars, N.cur_stream()),
                "ttd_refresh_copies")

    def apply_update(self, bump_global_step: bool = True, pdl: bool = False) -> None:
        """The fused push: reduce (all replicas) -> L2 -> SGD -> EMA -> refresh copies (all replicas)."""
        ema_ptr = self.ema.data_ptr() if self.hyper.ema_decay > 0 else None
        N.check(self.lib.ttd_apply_update(self.w.data_ptr(), ema_ptr, self.vars, self.nvars, self.lo, self.hi,
                                          C.byref(self._hyper_native), C.byref(self._peers), self.state_ptr,
                                          (1 if bump_global_step else 0) | (2 if pdl else 0), N.cur_stream()),
                "ttd_apply_update")

    def async_pull(self) -> None:
        """M1 of the async step: fetch every owner's master over NVLink into the local compute copies."""
        ema_ptr = self.ema.data_ptr() if self.hyper.ema_decay > 0 else None
        N.check(self.lib.ttd_async_pull(self.w.data_ptr(), ema_ptr, self.vars, self.nvars, self.lo, self.hi,
                                        C.byref(self._hyper_async), C.byref(self._async_peers), self.state_ptr,
                                        N.cur_stream()), "ttd_async_pull")

    def async_push(self) -> None:
        """M2 of the async step: apply this worker's gradient to the owners with NVLink atomics, bump global_step."""
        N.check(self.lib.ttd_async_push(self.w.data_ptr(), self.grads.data_ptr(), self.vars, self.nvars,
                                        C.byref(self._hyper_async), C.byref(self._async_peers), self.state_ptr,
                                        N.cur_stream()), "ttd_async_push")

    def snapshot_owner_shards(self) -> None:
        """Checkpoint path, ONE process: copy every owner's master / EMA range into the local buffers through the
        peer mapping (no collective, so the chief's 60-second timer needs nobody's cooperation)."""
        if self.comm is None:
            return
        base = self._arena.peer_ptrs
        for owner in range(self.world):
            lo, hi = self.plan.range_of(owner)
            if owner == self.rank or hi <= lo:
                continue
            for key, t in (("w", self.w), ("ema", self.ema)):
                src = base[owner] + self._off[key] + lo * 4
                N.check(self.lib.ttd_copy_d2d(t.data_ptr() + lo * 4, src, (hi - lo) * 4, N.cur_stream()), "ttd_copy_d2d")

    def zero_flags(self) -> None:
        self._raw[self._flags_off:self._flags_off + self._flags_bytes].zero_()

    # ---- host access (log cadence / checkpoints only) ---------------------------------------------
    def read_state(self) -> N.StepState:
        host = self.state.cpu().numpy().tobytes()
        return N.StepState.from_buffer_copy(host)

    def set_global_step(self, gs: int) -> None:
        self.state[0] = int(gs)

    def export_variables(self, with_ema: bool = True) -> Dict[str, np.ndarray]:
        """name -> array under the reference's variable names, kernels in TF layout."""
        w = self.w.cpu().numpy()
        out: Dict[str, np.ndarray] = {GLOBAL_STEP: np.asarray(int(self.state[0].item()), dtype=np.int64)}
        ema = self.ema.cpu().numpy() if (with_ema and self.hyper.ema_decay > 0) else None
        for name, (off, n, shape) in self.layout.slices().items():
            out[name] = w[off:off + n].reshape(shape).copy()
            if ema is not None:
                out[name + EMA_SUFFIX] = ema[off:off + n].reshape(shape).copy()
        return out

    def import_variables(self, tensors: Dict[str, np.ndarray]) -> None:
        w = self.w.cpu().numpy()
        ema = self.ema.cpu().numpy()
        for name, (off, n, shape) in self.layout.slices().items():
            if name not in tensors:
                raise KeyError(f"checkpoint lacks variable {name!r}")
            arr = np.asarray(tensors[name], dtype=np.float32)
            if tuple(arr.shape) != tuple(shape):
                raise ValueError(f"{name}: checkpoint shape {arr.shape} != model shape {shape}")
            w[off:off + n] = arr.reshape(-1)
            shadow = tensors.get(name + EMA_SUFFIX)
            ema[off:off + n] = (np.asarray(shadow, dtype=np.float32) if shadow is not None else arr).reshape(-1)
        self.w.copy_(torch.from_numpy(w))
        self.ema.copy_(torch.from_numpy(ema))
        if GLOBAL_STEP in tensors:
            self.set_global_step(int(np.asarray(tensors[GLOBAL_STEP])))
        self.refresh_copies()
