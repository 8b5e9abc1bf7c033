"""TF1 tensor-bundle checkpoints through ``libttd_io.so`` (C++: ``csrc/io/bundle.cc``, ``sstable.cc``).

This is the Saver of the framework: what ``MonitoredTrainingSession(checkpoint_dir=...)`` writes for the
reference (distribute_training.py:209-213) — ``model.ckpt-<step>.index`` + ``.data-0000k-of-0000N`` — under
the reference's variable names, with tensor *k*'s bytes in the shard of the ps task that owns it.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Mapping, Optional

import numpy as np

from .. import _build

DT_FLOAT, DT_DOUBLE, DT_INT32, DT_UINT8, DT_INT64, DT_BOOL, DT_BFLOAT16, DT_HALF = 1, 2, 3, 4, 9, 10, 14, 19
_NP_TO_DT = {np.dtype("float32"): DT_FLOAT, np.dtype("float64"): DT_DOUBLE, np.dtype("int32"): DT_INT32,
             np.dtype("uint8"): DT_UINT8, np.dtype("int64"): DT_INT64, np.dtype("bool"): DT_BOOL,
             np.dtype("float16"): DT_HALF}
_DT_TO_NP = {v: k for k, v in _NP_TO_DT.items()}
_DT_TO_NP[DT_BFLOAT16] = np.dtype("uint16")  # numpy has no bfloat16: raw bits

_lib = None


class BundleError(RuntimeError):
    pass


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        try:
            path = _build.build_io()
        except Exception as exc:  # noqa: BLE001
            path = _build.IO_LIB
            if not path.exists():
                raise BundleError(f"libttd_io.so is missing and could not be built: {exc}") from exc
        L = C.CDLL(str(path))
        L.ttdio_last_error.restype = C.c_char_p
        L.ttdio_crc32c.restype = C.c_uint32
        L.ttdio_crc32c.argtypes = [C.c_void_p, C.c_uint64]
        L.ttdio_crc32c_software.restype = C.c_uint32
        L.ttdio_crc32c_software.argtypes = [C.c_void_p, C.c_uint64]
        L.ttdio_crc32c_masked.restype = C.c_uint32
        L.ttdio_crc32c_masked.argtypes = [C.c_void_p, C.c_uint64]
        L.ttdio_bundle_write_shard.restype = C.c_int
        L.ttdio_bundle_write_shard.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_char_p),
                                               C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                               C.POINTER(C.POINTER(C.c_int64)), C.POINTER(C.c_void_p),
                                               C.POINTER(C.c_int64)]
        L.ttdio_bundle_merge.restype = C.c_int
        L.ttdio_bundle_merge.argtypes = [C.c_char_p, C.c_int]
        L.ttdio_bundle_open.restype = C.c_void_p
        L.ttdio_bundle_open.argtypes = [C.c_char_p]
        L.ttdio_bundle_close.argtypes = [C.c_void_p]
        L.ttdio_bundle_num_entries.argtypes = [C.c_void_p]
        L.ttdio_bundle_num_shards.argtypes = [C.c_void_p]
        L.ttdio_bundle_entry.restype = C.c_int
        L.ttdio_bundle_entry.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int32),
                                         C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_int32),
                                         C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_uint32)]
        L.ttdio_bundle_read.restype = C.c_int
        L.ttdio_bundle_read.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int64]
        L.ttdio_events_open.restype = C.c_void_p
        L.ttdio_events_open.argtypes = [C.c_char_p, C.c_double, C.c_char_p]
        L.ttdio_events_path.restype = C.c_char_p
        L.ttdio_events_path.argtypes = [C.c_void_p]
        L.ttdio_events_add_scalar.restype = C.c_int
        L.ttdio_events_add_scalar.argtypes = [C.c_void_p, C.c_char_p, C.c_float, C.c_int64, C.c_double]
        L.ttdio_events_flush.argtypes = [C.c_void_p]
        L.ttdio_events_close.argtypes = [C.c_void_p]
        L.ttdio_loader_create.restype = C.c_void_p
        L.ttdio_loader_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                          C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.ttdio_loader_push_epoch.restype = C.c_int
        L.ttdio_loader_push_epoch.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.ttdio_loader_backlog.restype = C.c_int64
        L.ttdio_loader_backlog.argtypes = [C.c_void_p]
        L.ttdio_loader_next.restype = C.c_int
        L.ttdio_loader_next.argtypes = [C.c_void_p, C.c_int32]
        L.ttdio_loader_release.restype = C.c_int
        L.ttdio_loader_release.argtypes = [C.c_void_p, C.c_int64]
        L.ttdio_loader_destroy.restype = None
        L.ttdio_loader_destroy.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def _err() -> str:
    return lib().ttdio_last_error().decode(errors="replace")


def crc32c(data: bytes) -> int:
    return lib().ttdio_crc32c(data, len(data))


def crc32c_software(data: bytes) -> int:
    return lib().ttdio_crc32c_software(data, len(data))


def masked_crc32c(data: bytes) -> int:
    return lib().ttdio_crc32c_masked(data, len(data))


def write_shard(prefix: str, shard_id: int, num_shards: int, tensors: Mapping[str, np.ndarray]) -> None:
    """Phase 1: write ``<prefix>.data-<shard_id>-of-<num_shards>`` and the partial index of this shard."""
    names = sorted(tensors)
    n = len(names)
    arrs = [np.asarray(tensors[k], order="C") for k in names]  # (ascontiguousarray would turn 0-d into 1-d)
    for k, a in zip(names, arrs):
        if a.dtype not in _NP_TO_DT:
            raise BundleError(f"{k}: dtype {a.dtype} has no TF DataType mapping")
    c_names = (C.c_char_p * n)(*[k.encode() for k in names])
    c_dt = (C.c_int32 * n)(*[_NP_TO_DT[a.dtype] for a in arrs])
    c_nd = (C.c_int32 * n)(*[a.ndim for a in arrs])
    shape_bufs = [(C.c_int64 * max(a.ndim, 1))(*a.shape) for a in arrs]
    c_shapes = (C.POINTER(C.c_int64) * n)(*[C.cast(b, C.POINTER(C.c_int64)) for b in shape_bufs])
    c_data = (C.c_void_p * n)(*[a.ctypes.data if a.size else None for a in arrs])
    c_nb = (C.c_int64 * n)(*[a.nbytes for a in arrs])
    rc = lib().ttdio_bundle_write_shard(prefix.encode(), shard_id, num_shards, n, c_names, c_dt, c_nd, c_shapes,
                                        c_data, c_nb)
    if rc != 0:
        raise BundleError(f"write_shard({prefix}, {shard_id}/{num_shards}) failed: {_err()}")


def merge(prefix: str, num_shards: int) -> None:
    """Phase 2 (chief): join the partial indices into ``<prefix>.index``."""
    if lib().ttdio_bundle_merge(prefix.encode(), num_shards) != 0:
        raise BundleError(f"merge({prefix}) failed: {_err()}")


def write_bundle(prefix: str, tensors: Mapping[str, np.ndarray], shard_of: Optional[Mapping[str, int]] = None,
                 num_shards: int = 1) -> None:
    """Single-process convenience: write every shard, then merge."""
    shard_of = shard_of or {}
    for k in range(num_shards):
        part = {name: t for name, t in tensors.items() if shard_of.get(name, 0) % num_shards == k}
        write_shard(prefix, k, num_shards, part)
    merge(prefix, num_shards)


class BundleReader:
    def __init__(self, prefix: str):
        self._h = lib().ttdio_bundle_open(prefix.encode())
        if not self._h:
            raise BundleError(f"cannot open bundle {prefix}: {_err()}")
        self.prefix = prefix
        self.entries: Dict[str, dict] = {}
        L = lib()
        for i in range(L.ttdio_bundle_num_entries(self._h)):
            name, dt, nd = C.c_char_p(), C.c_int32(), C.c_int32()
            shape, shard = (C.c_int64 * 8)(), C.c_int32()
            off, size, crc = C.c_int64(), C.c_int64(), C.c_uint32()
            if L.ttdio_bundle_entry(self._h, i, C.byref(name), C.byref(dt), C.byref(nd), shape, C.byref(shard),
                                    C.byref(off), C.byref(size), C.byref(crc)) != 0:
                raise BundleError(_err())
            self.entries[name.value.decode()] = dict(dtype=dt.value, shape=tuple(shape[:nd.value]), shard_id=shard.value,
                                                    offset=off.value, size=size.value, crc32c=crc.value)

    @property
    def num_shards(self) -> int:
        return lib().ttdio_bundle_num_shards(self._h)

    def keys(self) -> List[str]:
        return sorted(self.entries)

    def read(self, name: str) -> np.ndarray:
        e = self.entries.get(name)
        if e is None:
            raise KeyError(name)
        dt = _DT_TO_NP.get(e["dtype"])
        if dt is None:
            raise BundleError(f"{name}: unsupported TF dtype code {e['dtype']}")
        out = np.empty(e["shape"], dtype=dt)
        if out.nbytes != e["size"]:
            raise BundleError(f"{name}: shape {e['shape']} x {dt} != stored size {e['size']}")
        buf = out.ctypes.data if out.size else None
        if lib().ttdio_bundle_read(self._h, name.encode(), buf, e["size"]) != 0:
            raise BundleError(f"read({name}) failed: {_err()}")
        return out

    def read_all(self) -> Dict[str, np.ndarray]:
        return {k: self.read(k) for k in self.keys()}

    def close(self) -> None:
        if self._h:
            lib().ttdio_bundle_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def read_bundle(prefix: str) -> Dict[str, np.ndarray]:
    with BundleReader(prefix) as r:
        return r.read_all()
