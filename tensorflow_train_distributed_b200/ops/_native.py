"""ctypes binding of ``libttd_b200.so`` (declarations in ``include/ttd_b200.h``).

The library is built in-tree by :mod:`tensorflow_train_distributed_b200._build`. On a machine with a
GPU a missing library is an error, never a silent fallback to PyTorch ops: the kernels *are* the
product.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import torch  # noqa: F401  (loads libcudart.so.12 before our library needs it)

from .. import _build

MAX_PEERS = 8

EPI_FWD, EPI_DGRAD, EPI_WGRAD, EPI_FWD_XENT, EPI_WGRAD_UPDATE = 0, 1, 2, 3, 4
ACT_NONE, ACT_RELU, ACT_ELU = 0, 1, 2
ACT_CODES = {"none": ACT_NONE, "relu": ACT_RELU, "elu": ACT_ELU}
REG_CODES = {"none": 0, "l1": 1, "l2": 2}


class GemmDesc(C.Structure):
    _fields_ = [
        ("a", C.c_void_p), ("lda", C.c_int64),
        ("b", C.c_void_p), ("ldb", C.c_int64),
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("block_n", C.c_int32),
        ("mode", C.c_int32),
        ("act", C.c_int32),
        ("drop_rate", C.c_float),
        ("layer_id", C.c_uint32),
        ("bias", C.c_void_p),
        ("out", C.c_void_p), ("ldo", C.c_int64),
        ("outT", C.c_void_p), ("ldt", C.c_int64),
        ("outf", C.c_void_p), ("ldf", C.c_int64),
        ("aux", C.c_void_p), ("ldaux", C.c_int64),
        ("scale", C.c_float),
        ("labels", C.c_void_p),
        ("metrics", C.c_void_p),
        ("b_peer", C.c_void_p * 8),
        ("k_bound", C.c_int32 * 9),
        ("n_b_peers", C.c_int32),
        ("upd_regularize", C.c_int32),
        ("upd_w", C.c_void_p),
        ("upd_ema", C.c_void_p),
        ("upd_bias", C.c_void_p),
        ("a_slot_rows", C.c_int32), ("a_slots", C.c_int32),
    ]


class VarDesc(C.Structure):
    _fields_ = [
        ("offset", C.c_int64),
        ("rows", C.c_int32), ("cols", C.c_int32),
        ("regularize", C.c_int32), ("reserved", C.c_int32),
        ("w_rm", C.c_void_p), ("ld_rm", C.c_int64),
        ("w_t", C.c_void_p), ("ld_t", C.c_int64),
        ("bias_c", C.c_void_p),
        ("w_f32", C.c_void_p), ("ld_f32", C.c_int64),
        ("wt_f32", C.c_void_p), ("ld_tf32", C.c_int64),
    ]


class Hyper(C.Structure):
    _fields_ = [
        ("lr0", C.c_float),
        ("decay_rate", C.c_float),
        ("decay_steps", C.c_int64),
        ("staircase", C.c_int32),
        ("reg_kind", C.c_int32),
        ("reg_scale", C.c_float),
        ("ema_decay", C.c_float),
        ("ema_warmup", C.c_int32),
        ("grad_scale", C.c_float),
        ("host_mirror", C.c_void_p),
        ("mirror_slots", C.c_int32),
        ("reserved", C.c_int32),
    ]


class StepState(C.Structure):
    """Host mirror of the 64-byte device step-state block."""
    _fields_ = [
        ("global_step", C.c_int64),
        ("local_step", C.c_uint64),
        ("metrics_acc", C.c_float * 2),
        ("metrics_out", C.c_float * 2),
        ("done_counter", C.c_uint32),
        ("lr_last", C.c_float),
        ("reserved", C.c_uint32 * 6),
    ]


STATE_BYTES = C.sizeof(StepState)
assert STATE_BYTES == 64, STATE_BYTES
# byte offsets inside the state block
OFF_GLOBAL_STEP = StepState.global_step.offset
OFF_LOCAL_STEP = StepState.local_step.offset
OFF_METRICS_ACC = StepState.metrics_acc.offset
OFF_METRICS_OUT = StepState.metrics_out.offset


class Peers(C.Structure):
    _fields_ = [
        ("grads", C.c_void_p * MAX_PEERS),
        ("copy_delta", C.c_int64 * MAX_PEERS),
        ("flags", C.c_void_p * MAX_PEERS),
        ("grads_mc", C.c_void_p),
        ("mc_delta", C.c_int64),
        ("n_grads", C.c_int32), ("n_copies", C.c_int32), ("rank", C.c_int32),
        ("barrier", C.c_int32),
        ("grid", C.c_int32),
        ("reserved", C.c_int32),
    ]


class AsyncPeers(C.Structure):
    _fields_ = [
        ("w_peer", C.c_void_p * MAX_PEERS),
        ("bounds", C.c_int64 * (MAX_PEERS + 1)),
        ("state0", C.c_void_p),
        ("n_owners", C.c_int32), ("rank", C.c_int32),
    ]


class ConvDesc(C.Structure):
    _fields_ = [("B", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("Cin", C.c_int32), ("Cout", C.c_int32),
                ("KH", C.c_int32), ("KW", C.c_int32), ("same_pad", C.c_int32)]


TAIL_MAX_LAYERS = 6


class TailLayer(C.Structure):
    _fields_ = [
        ("in_", C.c_int32), ("out", C.c_int32), ("act_kind", C.c_int32), ("drop_rate", C.c_float),
        ("layer_id", C.c_uint32), ("reserved", C.c_int32),
        ("w", C.c_void_p), ("ld_w", C.c_int64), ("wt", C.c_void_p), ("ld_wt", C.c_int64),
        ("bias", C.c_void_p), ("dw", C.c_void_p),
        ("act", C.c_void_p), ("ld_act", C.c_int64), ("actT", C.c_void_p), ("ld_actT", C.c_int64),
        ("dz", C.c_void_p), ("ld_dz", C.c_int64),
    ]


class TailDesc(C.Structure):
    _fields_ = [
        ("B", C.c_int32), ("n_layers", C.c_int32), ("cluster", C.c_int32), ("in_act_kind", C.c_int32),
        ("in_drop_rate", C.c_float), ("in_layer_id", C.c_uint32),
        ("a_in", C.c_void_p), ("ld_a_in", C.c_int64), ("a_inT", C.c_void_p), ("ld_a_inT", C.c_int64),
        ("dz_in", C.c_void_p), ("ld_dz_in", C.c_int64), ("dz_inT", C.c_void_p), ("ld_dz_inT", C.c_int64),
        ("labels", C.c_void_p), ("metrics", C.c_void_p), ("step_ptr", C.c_void_p),
        ("layer", TailLayer * TAIL_MAX_LAYERS),
    ]


COL_MAX_LAYERS = 6


class ColumnLayer(C.Structure):
    _fields_ = [
        ("in_", C.c_int32), ("out", C.c_int32), ("layer_id", C.c_uint32), ("reserved", C.c_int32),
        ("w_t", C.c_void_p), ("ld_t", C.c_int64), ("w_rm", C.c_void_p), ("ld_rm", C.c_int64),
        ("bias", C.c_void_p),
        ("aT", C.c_void_p), ("ld_aT", C.c_int64), ("dzT", C.c_void_p), ("ld_dzT", C.c_int64),
    ]


class ColumnDesc(C.Structure):
    _fields_ = [
        ("B", C.c_int32), ("n_layers", C.c_int32), ("act", C.c_int32), ("drop_rate", C.c_float),
        ("cluster", C.c_int32), ("pool_batches", C.c_int32),
        ("x_pool", C.c_void_p), ("ld_x", C.c_int64),
        ("labels_pool", C.c_void_p), ("metrics", C.c_void_p), ("step_ptr", C.c_void_p),
        ("global_step_ptr", C.c_void_p), ("lr_out", C.c_void_p),
        ("hyper", Hyper),
        ("layer", ColumnLayer * COL_MAX_LAYERS),
    ]


class NativeError(RuntimeError):
    pass


_lib = None


def lib_path() -> Path:
    return _build.CUDA_LIB


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load (building first if the sources changed or the library is absent) the kernel library."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if build_if_missing:
        try:
            path = _build.build_cuda()
        except Exception as exc:  # no nvcc on this machine: fall back to a shipped binary
            if not path.exists():
                raise NativeError(f"libttd_b200.so is missing and could not be built: {exc}") from exc
    if not path.exists():
        raise NativeError(f"{path} not found; run `python -m tensorflow_train_distributed_b200._build`")
    lib = C.CDLL(str(path))
    lib.ttd_last_error.restype = C.c_char_p
    lib.ttd_version.restype = C.c_char_p
    lib.ttd_launch_count.restype = C.c_uint64
    lib.ttd_gemm_group.restype = C.c_int
    lib.ttd_gemm_group.argtypes = [C.POINTER(GemmDesc), C.c_int, C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p]
    lib.ttd_gemm_chain_problem_bytes.restype = C.c_uint64
    lib.ttd_gemm_chain_build.restype = C.c_int
    lib.ttd_gemm_chain_build.argtypes = [C.POINTER(GemmDesc), C.c_int, C.c_void_p, C.c_void_p]
    lib.ttd_gemm_chain_launch.restype = C.c_int
    lib.ttd_gemm_chain_launch.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p]
    lib.ttd_gemm_net_bytes.restype = C.c_uint64
    lib.ttd_gemm_net_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.ttd_gemm_net_build.restype = C.c_int
    lib.ttd_gemm_net_build.argtypes = [C.POINTER(GemmDesc), C.c_int, C.POINTER(C.c_uint32), C.c_int, C.c_int, C.c_void_p,
                                       C.c_void_p]
    lib.ttd_gemm_net_launch.restype = C.c_int
    lib.ttd_gemm_net_launch.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_uint32,
                                        C.c_void_p]
    lib.ttd_gemm_net_launch_ex.restype = C.c_int
    lib.ttd_gemm_net_launch_ex.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_void_p,
                                           C.c_uint32, C.POINTER(Hyper), C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]
    lib.ttd_mlp_column_create.restype = C.c_void_p
    lib.ttd_mlp_column_create.argtypes = [C.POINTER(ColumnDesc)]
    lib.ttd_mlp_column_destroy.restype = None
    lib.ttd_mlp_column_destroy.argtypes = [C.c_void_p]
    lib.ttd_mlp_column_launch.restype = C.c_int
    lib.ttd_mlp_column_launch.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p]
    lib.ttd_mlp_column_set_trace.restype = None
    lib.ttd_mlp_column_set_trace.argtypes = [C.c_void_p]
    lib.ttd_mlp_column_fits.restype = C.c_int
    lib.ttd_mlp_column_fits.argtypes = [C.c_int32, C.POINTER(C.c_int32), C.c_int32]
    lib.ttd_gemm_net_set_trace.restype = None
    lib.ttd_gemm_net_set_trace.argtypes = [C.c_void_p]
    lib.ttd_softmax_xent.restype = C.c_int
    lib.ttd_softmax_xent.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64,
                                     C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p]
    lib.ttd_prep_input.restype = C.c_int
    lib.ttd_prep_input.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                   C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ttd_prep_input_ex.restype = C.c_int
    lib.ttd_prep_input_ex.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                      C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
    lib.ttd_feeder_create.restype = C.c_void_p
    lib.ttd_feeder_create.argtypes = [C.c_void_p] * 6 + [C.c_size_t, C.c_size_t]
    lib.ttd_feeder_destroy.restype = None
    lib.ttd_feeder_destroy.argtypes = [C.c_void_p]
    lib.ttd_feeder_set_prep.restype = C.c_int
    lib.ttd_feeder_set_prep.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]
    lib.ttd_feeder_feed.restype = C.c_int
    lib.ttd_feeder_feed.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
    lib.ttd_feeder_before_step.restype = C.c_int
    lib.ttd_feeder_before_step.argtypes = [C.c_void_p]
    lib.ttd_feeder_after_step.restype = C.c_int
    lib.ttd_feeder_after_step.argtypes = [C.c_void_p]
    lib.ttd_feeder_step.restype = C.c_int
    lib.ttd_feeder_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    lib.ttd_init_params.restype = C.c_int
    lib.ttd_init_params.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(VarDesc), C.c_int, C.c_uint64, C.c_void_p]
    lib.ttd_refresh_copies.restype = C.c_int
    lib.ttd_refresh_copies.argtypes = [C.c_void_p, C.POINTER(VarDesc), C.c_int, C.c_void_p]
    lib.ttd_apply_update.restype = C.c_int
    lib.ttd_apply_update.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(VarDesc), C.c_int, C.c_int64, C.c_int64,
                                     C.POINTER(Hyper), C.POINTER(Peers), C.c_void_p, C.c_int32, C.c_void_p]
    lib.ttd_async_pull.restype = C.c_int
    lib.ttd_async_pull.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(VarDesc), C.c_int, C.c_int64, C.c_int64,
                                   C.POINTER(Hyper), C.POINTER(AsyncPeers), C.c_void_p, C.c_void_p]
    lib.ttd_async_push.restype = C.c_int
    lib.ttd_async_push.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(VarDesc), C.c_int, C.POINTER(Hyper),
                                   C.POINTER(AsyncPeers), C.c_void_p, C.c_void_p]
    lib.ttd_conv_pool_fwd.restype = C.c_int
    lib.ttd_conv_pool_fwd.argtypes = [C.POINTER(ConvDesc), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]
    lib.ttd_conv_pool_bwd.restype = C.c_int
    lib.ttd_conv_pool_bwd.argtypes = [C.POINTER(ConvDesc), C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                      C.c_int64, C.c_void_p]
    lib.ttd_mlp_tail.restype = C.c_int
    lib.ttd_mlp_tail.argtypes = [C.POINTER(TailDesc), C.c_uint64, C.c_uint32, C.c_void_p]
    lib.ttd_set_barrier_timeout_ms.restype = None
    lib.ttd_set_barrier_timeout_ms.argtypes = [C.c_uint64]
    lib.ttd_copy_d2d.restype = C.c_int
    lib.ttd_copy_d2d.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().ttd_last_error().decode(errors="replace")
        raise NativeError(f"{what} failed (code {rc}): {msg}")


def launch_count() -> int:
    return int(load().ttd_launch_count()) if _lib is not None else 0


def ptr(t) -> int | None:
    """Device address of a tensor (None stays NULL)."""
    return None if t is None else t.data_ptr()


def cur_stream() -> int:
    return torch.cuda.current_stream().cuda_stream
