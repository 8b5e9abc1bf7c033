"""Native TCP parameter server (``--transport tcp``): the ps task is ``csrc/io/ps_server.cc`` listening on its own
``--ps_hosts`` entry, the way the reference's ``tf.train.Server`` does (distribute_training.py:173-181); workers
connect to every ps directly. Same protocol and semantics as the gloo tier (:mod:`.cpu_ps`), but no rendezvous
process group: a worker that dies just drops its connections and a restarted one reconnects, the ps keeps its state.
"""
from __future__ import annotations

import ctypes as C
import socket
import struct
import time
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from ..cluster import ClusterSpec, TaskInfo
from ..config import TrainConfig
from ..session import TransportError
from .cpu_ps import make_placement
from .sharding import GLOBAL_STEP, ParamLayout

OP_HELLO = 9
REG_CODES = {"none": 0, "l1": 1, "l2": 2}


class PsConfig(C.Structure):
    _fields_ = [("numel", C.c_int64), ("n_workers", C.c_int32), ("n_agg", C.c_int32), ("tokens_per_step", C.c_int32),
                ("owns_global_step", C.c_int32), ("reg_kind", C.c_int32), ("use_ema", C.c_int32),
                ("reg_scale", C.c_float), ("ema_decay", C.c_float), ("idle_timeout_secs", C.c_double)]


def _lib():
    from ..io.bundle import lib
    L = lib()
    if not getattr(L, "_ttdps_bound", False):
        L.ttdps_server_create.restype = C.c_void_p
        L.ttdps_server_create.argtypes = [C.c_char_p, C.c_int32, C.c_void_p, C.c_void_p]
        L.ttdps_server_port.restype = C.c_int32
        L.ttdps_server_port.argtypes = [C.c_void_p]
        L.ttdps_server_serve.restype = C.c_int
        L.ttdps_server_serve.argtypes = [C.c_void_p]
        L.ttdps_server_stats.restype = None
        L.ttdps_server_stats.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        L.ttdps_server_destroy.restype = None
        L.ttdps_server_destroy.argtypes = [C.c_void_p]
        L._ttdps_bound = True
    return L


def _split(hostport: str) -> Tuple[str, int]:
    host, _, port = hostport.rpartition(":")
    return (host or "127.0.0.1"), int(port)


class NativePsServer:
    """``server.join()`` (:179-181) in C++: returns once every worker has said goodbye (or the idle limit passes)."""

    def __init__(self, cfg: TrainConfig, cluster: ClusterSpec, task: TaskInfo, layout: ParamLayout):
        self.cfg, self.ps_index = cfg, task.index
        place = make_placement(layout, cluster.num_tasks("ps"))
        specs = {v.name: v for v in place.variables}
        names = place.names_on(task.index)
        numel = place.numel_on(task.index)
        mask = np.zeros(max(numel, 1), dtype=np.float32)
        off = 0
        for n in names:
            k = specs[n].numel
            if specs[n].regularize:
                mask[off:off + k] = 1.0
            off += k
        n_workers = cluster.num_tasks("worker")
        pc = PsConfig(numel, n_workers, cfg.aggregate_n, max(cfg.total_num_replicas or n_workers, cfg.aggregate_n),
                      1 if place.owner[GLOBAL_STEP] == task.index else 0, REG_CODES[cfg.regularizer],
                      1 if cfg.use_moving_average else 0, cfg.regularizer_scale, cfg.moving_average_decay,
                      float(cfg.ps_idle_timeout_secs))
        host, port = _split(cluster.job_tasks("ps")[task.index])
        self._lib = _lib()
        self._h = self._lib.ttdps_server_create(host.encode(), port, C.byref(pc), mask.ctypes.data)
        if not self._h:
            raise RuntimeError("ps: " + self._lib.ttdio_last_error().decode(errors="replace"))
        self.port = self._lib.ttdps_server_port(self._h)

    def serve(self) -> Dict[str, int]:
        rc = self._lib.ttdps_server_serve(self._h)
        out = (C.c_int64 * 8)()
        self._lib.ttdps_server_stats(self._h, out)
        stats = {"applied": int(out[0]), "dropped_stale": int(out[1]), "pulls": int(out[2]), "global_step": int(out[3])}
        if rc == 1:
            stats["idle_timeout"] = 1
        return stats

    def close(self) -> None:
        if self._h:
            self._lib.ttdps_server_destroy(self._h)
            self._h = None


class TcpChannel:
    """Request/reply connections from one worker to every ps (header = 4 x int64, payload = float32)."""

    def __init__(self, cluster: ClusterSpec, worker_index: int, timeout_s: float = 120.0):
        self.addrs = [_split(h) for h in cluster.job_tasks("ps")]
        self.worker_index, self.timeout_s = worker_index, timeout_s
        self.socks: Dict[int, socket.socket] = {}

    def _sock(self, ps: int) -> socket.socket:
        s = self.socks.get(ps)
        if s is not None:
            return s
        deadline = time.time() + self.timeout_s
        last: Optional[Exception] = None
        while time.time() < deadline:  # the ps may still be starting (the reference's workers retry the same way)
            try:
                s = socket.create_connection(self.addrs[ps], timeout=5.0)
                s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                s.settimeout(None)
                self.socks[ps] = s
                self.call(ps, OP_HELLO, self.worker_index)
                return s
            except OSError as exc:
                last = exc
                self.socks.pop(ps, None)
                time.sleep(0.1)
        raise TransportError(f"cannot reach parameter server {ps} at {self.addrs[ps]}: {last}")

    @staticmethod
    def _recv_into(s: socket.socket, view: memoryview) -> None:
        got = 0
        while got < len(view):
            r = s.recv_into(view[got:])
            if r == 0:
                raise ConnectionError("connection closed by the parameter server")
            got += r

    def call(self, ps: int, op: int, a: int = 0, b: int = 0, c: int = 0, payload: Optional[torch.Tensor] = None,
             reply_numel: int = 0):
        try:
            s = self._sock(ps)
            s.sendall(struct.pack("<4q", op, a, b, c))
            if payload is not None and payload.numel():
                s.sendall(memoryview(payload.contiguous().numpy()).cast("B"))
            raw = bytearray(32)
            self._recv_into(s, memoryview(raw))
            hdr = torch.tensor(struct.unpack("<4q", raw), dtype=torch.int64)
            body = None
            if reply_numel:
                body = torch.empty(reply_numel, dtype=torch.float32)
                self._recv_into(s, memoryview(body.numpy()).cast("B"))
            return hdr, body
        except (OSError, ConnectionError) as exc:
            self.drop(ps)
            raise TransportError(f"lost parameter server {ps}: {exc}") from exc

    def drop(self, ps: Optional[int] = None) -> None:
        for k in ([ps] if ps is not None else list(self.socks)):
            s = self.socks.pop(k, None)
            if s is not None:
                try:
                    s.close()
                except OSError:
                    pass

    def close(self) -> None:
        self.drop()
