"""GPU parameter store: the "ps" of the reference as device memory.

What lives here replaces the ps job of the reference (``server.join()``, distribute_training.py:179-181, and
the ``ApplyGradientDescent`` / ``AssignAdd(global_step)`` ops TF places on it, :150-152):

* ``w``      flat fp32 master parameters, one ``[rows+1, cols]`` group per layer (kernel + bias)
* ``ema``    flat fp32 shadow variables (``<var>/ExponentialMovingAverage``)
* ``grads``  flat fp32 gradient buffer — the wgrad GEMMs write straight into it; on several GPUs it lives in
             symmetric memory so the owner of a range can read every replica's copy over NVLink
* ``copies`` bf16 compute copies in the two K-major layouts the GEMMs need (``w_t`` for forward, ``w_rm``
             for dgrad); refreshed by the update kernel on *every* replica (that is the parameter pull)
* ``state``  64-byte device block: global_step, local_step, metric accumulators (never read by the host on the
             step path)
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from ..ops import _native as N
from .sharding import EMA_SUFFIX, GLOBAL_STEP, FlatPlan, ParamLayout, flat


def _pad8(n: int) -> int:
    return -(-n // 8) * 8


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
                 need_w_rm_first: bool = False):
        """``comm`` is a :class:`~.symmetric.SymmetricComm` for multi-GPU runs (None: single GPU)."""
        self.layout, self.hyper, self.device, self.comm = layout, hyper, device, comm
        self.lib = N.load()
        self.world = comm.world_size if comm is not None else 1
        self.rank = comm.rank if comm is not None else 0
        total = layout.total
        self.w = torch.zeros(total, device=device, dtype=torch.float32)
        self.ema = torch.zeros(total, device=device, dtype=torch.float32)
        self.state = torch.zeros(8, device=device, dtype=torch.int64)  # TtdStepState (64 B)

        # ---- bf16 compute copies: one arena, carved per layer -----------------------------------
        sizes = []
        for i, g in enumerate(layout.groups):
            want_rm = need_w_rm_first or i > 0  # dgrad of the first layer is never needed (input has no grad)
            rm = g.rows * _pad8(g.cols) if want_rm else 0
            t = g.cols * _pad8(g.rows)
            sizes.append((rm, t))
        arena_elems = sum(-(-a // 128) * 128 + -(-b // 128) * 128 for a, b in sizes)
        grads_bytes = total * 4
        copies_bytes = arena_elems * 2
        if comm is not None:
            # grads | copies | flags   in one symmetric allocation, so peer addresses differ by one constant
            self._flags_bytes = 148 * N.MAX_PEERS * 4
            self._arena = comm.alloc(grads_bytes + copies_bytes + self._flags_bytes)
            raw = self._arena.local  # uint8 tensor
            raw.zero_()
            self.grads = raw[:grads_bytes].view(torch.float32)
            self.copies = raw[grads_bytes:grads_bytes + copies_bytes].view(torch.bfloat16)
            self._flags_off = grads_bytes + copies_bytes
        else:
            self._arena = None
            self.grads = torch.zeros(total, device=device, dtype=torch.float32)
            self.copies = torch.zeros(arena_elems, device=device, dtype=torch.bfloat16)
        self.w_rm: List[Optional[torch.Tensor]] = []
        self.w_t: List[torch.Tensor] = []
        off = 0
        for g, (rm, t) in zip(layout.groups, sizes):
            if rm:
                self.w_rm.append(self.copies[off:off + rm].view(g.rows, _pad8(g.cols)))
                off += -(-rm // 128) * 128
            else:
                self.w_rm.append(None)
            self.w_t.append(self.copies[off:off + t].view(g.cols, _pad8(g.rows)))
            off += -(-t // 128) * 128

        # ---- variable table ----------------------------------------------------------------------
        descs = []
        for g, rm, t in zip(layout.groups, self.w_rm, self.w_t):
            v = N.VarDesc()
            v.offset, v.rows, v.cols = g.offset, g.rows, g.cols
            v.regularize = 1 if g.regularize else 0
            if rm is not None:
                v.w_rm, v.ld_rm = rm.data_ptr(), rm.stride(0)
            v.w_t, v.ld_t = t.data_ptr(), t.stride(0)
            descs.append(v)
        self.vars = (N.VarDesc * len(descs))(*descs)
        self.nvars = len(descs)
        self.plan: FlatPlan = flat(layout, self.world)
        self.lo, self.hi = self.plan.range_of(self.rank)
        self._peers = self._build_peers()
        self._hyper_native = hyper.to_native()

    # ------------------------------------------------------------------------------------------
    def _build_peers(self) -> N.Peers:
        p = N.Peers()
        if self.comm is None:
            p.grads[0] = self.grads.data_ptr()
            p.n_grads = p.n_copies = 1
            return p
        base = self._arena.peer_ptrs
        for r in range(self.world):
            p.grads[r] = base[r]
            p.copy_delta[r] = base[r] - base[self.rank]
            p.flags[r] = base[r] + self._flags_off
        p.grads_mc = None
        p.n_grads = p.n_copies = self.world
        p.rank = self.rank
        p.barrier = 1
        p.grid = min(148, max(1, -(-(self.hi - self.lo) // (32 * 32 * 8))))
        return p

    # ---- views ---------------------------------------------------------------------------------
    def bias(self, i: int) -> torch.Tensor:
        g = self.layout.groups[i]
        o = g.offset + g.rows * g.cols
        return self.w[o:o + g.cols]

    def grad_matrix(self, i: int) -> torch.Tensor:
        g = self.layout.groups[i]
        return self.grads[g.offset:g.offset + g.numel].view(g.rows + 1, g.cols)

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

    def apply_update(self, bump_global_step: bool = True) -> None:
        """The fused push: reduce (all replicas) -> L2 -> SGD -> EMA -> refresh copies (all replicas)."""
        ema_ptr = self.ema.data_ptr() if self.hyper.ema_decay > 0 else None
        N.check(self.lib.ttd_apply_update(self.w.data_ptr(), ema_ptr, self.vars, self.nvars, self.lo, self.hi,
                                          C.byref(self._hyper_native), C.byref(self._peers), self.state_ptr,
                                          1 if bump_global_step else 0, N.cur_stream()), "ttd_apply_update")

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
