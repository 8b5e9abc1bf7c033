"""The C++ parameter-server task (csrc/io/ps_server.cc) driven request by request: exact SyncReplicasOptimizer
arithmetic — mean of N fresh gradients, ONE global_step per aggregate, stale gradients dropped, token
queue — and the async apply with L2 + EMA, checked against closed forms."""
import ctypes as C
import struct
import threading

import numpy as np
import pytest
import torch

from tensorflow_train_distributed_b200.cluster import ClusterSpec
from tensorflow_train_distributed_b200.parallel import tcp_ps
from tensorflow_train_distributed_b200.parallel.cpu_ps import (OP_DONE, OP_GET_STATE, OP_INIT, OP_PULL, OP_PUSH_ASYNC,
                                                               OP_PUSH_SYNC, OP_READY, OP_SYNC_INIT, OP_TOKEN)
from tensorflow_train_distributed_b200.utils.schedule import ema_decay


class _Ps:
    def __init__(self, numel, n_workers=2, n_agg=2, reg_kind=0, reg_scale=0.0, use_ema=1, mask=None, idle=0.0):
        self.lib = tcp_ps._lib()
        cfg = tcp_ps.PsConfig(numel, n_workers, n_agg, max(n_workers, n_agg), 1, reg_kind, use_ema, reg_scale, 0.99, idle)
        m = np.ascontiguousarray(mask if mask is not None else np.zeros(numel), dtype=np.float32)
        self.h = self.lib.ttdps_server_create(b"127.0.0.1", 0, C.byref(cfg), m.ctypes.data)
        assert self.h, self.lib.ttdio_last_error()
        self.port = self.lib.ttdps_server_port(self.h)
        self.rc = None
        self.thread = threading.Thread(target=self._serve, daemon=True)
        self.thread.start()
        self.cluster = ClusterSpec({"ps": [f"127.0.0.1:{self.port}"], "worker": ["127.0.0.1:1", "127.0.0.1:2"]})

    def _serve(self):
        self.rc = self.lib.ttdps_server_serve(self.h)

    def channel(self, worker):
        return tcp_ps.TcpChannel(self.cluster, worker, timeout_s=10.0)

    def stats(self):
        out = (C.c_int64 * 8)()
        self.lib.ttdps_server_stats(self.h, out)
        return dict(applied=out[0], dropped=out[1], pulls=out[2], gs=out[3], init=out[4], done=out[5])

    def close(self):
        self.thread.join(timeout=10)
        self.lib.ttdps_server_destroy(self.h)


def _lr_bits(lr):
    return struct.unpack("<q", struct.pack("<d", lr))[0]


def test_sync_accumulator_tokens_and_stale_drop():
    ps = _Ps(numel=4)
    a, b = ps.channel(0), ps.channel(1)
    hdr, _ = a.call(0, OP_READY)
    assert int(hdr[1]) == 0  # not initialised yet: non-chief workers keep polling (wait_for_session)
    w0 = torch.tensor([1.0, 2.0, 3.0, 4.0])
    a.call(0, OP_INIT, 10, payload=torch.cat([w0, w0]))  # chief restores at global_step 10
    hdr, _ = b.call(0, OP_READY)
    assert int(hdr[1]) == 1 and int(hdr[2]) == 10
    a.call(0, OP_SYNC_INIT, 2)  # token queue pre-loaded with replicas_to_aggregate tokens of the current step
    lr = 0.5
    g1, g2 = torch.tensor([1.0, 0.0, 2.0, 4.0]), torch.tensor([3.0, 2.0, 0.0, 0.0])
    hdr, _ = a.call(0, OP_PUSH_SYNC, _lr_bits(lr), 10, payload=g1)
    assert int(hdr[1]) == 10 and ps.stats()["applied"] == 0  # one fresh gradient: nothing applied yet
    hdr, _ = b.call(0, OP_PUSH_SYNC, _lr_bits(lr), 10, payload=g2)
    assert int(hdr[1]) == 11  # the second one completes the aggregate: ONE global step
    hdr, w = a.call(0, OP_PULL, reply_numel=4)
    assert torch.allclose(w, w0 - lr * (g1 + g2) / 2) and int(hdr[1]) == 11
    # tokens: two pre-loaded (value 10) come out first, then the two of step 11
    toks = [int(a.call(0, OP_TOKEN)[0][1]), int(b.call(0, OP_TOKEN)[0][1]), int(a.call(0, OP_TOKEN)[0][1])]
    assert toks == [10, 10, 11]
    # a gradient tagged with the old local step is stale: dropped silently, nothing moves
    hdr, _ = a.call(0, OP_PUSH_SYNC, _lr_bits(lr), 10, payload=g1)
    assert int(hdr[1]) == 11 and ps.stats()["dropped"] == 1 and ps.stats()["applied"] == 1
    # EMA followed the apply with the warm-up decay min(0.99, (1+n)/(10+n)), n = new global step
    hdr, st = a.call(0, OP_GET_STATE, reply_numel=8)
    d = ema_decay(0.99, 11)
    assert torch.allclose(st[4:], w0 - (1 - d) * (w0 - st[:4]), atol=1e-6)
    # a worker blocked on an empty token queue is released when a peer leaves
    b.call(0, OP_TOKEN)  # the last token (11)
    got = {}
    t = threading.Thread(target=lambda: got.update(hdr=b.call(0, OP_TOKEN)[0]))
    t.start()
    t.join(timeout=0.3)
    assert t.is_alive()  # still blocked: queue empty
    a.call(0, OP_DONE)
    t.join(timeout=5)
    assert not t.is_alive() and int(got["hdr"][2]) == 11
    b.call(0, OP_DONE)
    ps.close()
    assert ps.rc == 0 and ps.stats()["done"] == 2


@pytest.mark.parametrize("reg_kind", [0, 1, 2])
def test_async_apply_with_regulariser_and_global_step(reg_kind):
    mask = np.array([1, 1, 0, 0], dtype=np.float32)  # hidden kernel vs bias / output layer (:55 vs :61)
    ps = _Ps(numel=4, n_workers=1, n_agg=1, reg_kind=reg_kind, reg_scale=0.1, use_ema=0, mask=mask)
    c = ps.channel(0)
    w0 = torch.tensor([1.0, -2.0, 3.0, -4.0])
    c.call(0, OP_INIT, 0, payload=torch.cat([w0, w0]))
    g = torch.tensor([0.5, 0.5, 0.5, 0.5])
    w = w0.clone()
    for step in range(3):
        hdr, _ = c.call(0, OP_PUSH_ASYNC, _lr_bits(0.1), step, payload=g)
        assert int(hdr[1]) == step + 1  # every apply bumps global_step (:152)
        m = torch.from_numpy(mask)
        reg = 0.1 * w * m if reg_kind == 2 else (0.1 * torch.sign(w) * m if reg_kind == 1 else 0.0)
        w = w - 0.1 * (g + reg)
    _, got = c.call(0, OP_PULL, reply_numel=4)
    assert torch.allclose(got, w, atol=1e-6)
    c.call(0, OP_DONE)
    ps.close()
    assert ps.stats()["applied"] == 3 and ps.stats()["pulls"] == 1


def test_idle_watchdog_ends_serve():
    ps = _Ps(numel=2, n_workers=1, n_agg=1, idle=0.3)
    c = ps.channel(0)
    c.call(0, OP_INIT, 0, payload=torch.zeros(4))
    ps.thread.join(timeout=5)
    assert ps.rc == 1  # nobody said goodbye, the idle limit ended serve()
    c.close()
    ps.lib.ttdps_server_destroy(ps.h)


def test_server_survives_garbage_and_keeps_serving():
    """Malformed headers, unknown ops, short and oversized payloads, abrupt closes: the ps neither crashes nor wedges."""
    import random
    import socket
    ps = _Ps(numel=64, n_workers=1, n_agg=1)
    rnd = random.Random(7)
    for _ in range(40):
        s = socket.create_connection(("127.0.0.1", ps.port))
        s.settimeout(0.05)
        try:
            for _ in range(rnd.randrange(1, 4)):
                op = rnd.choice([0, 1, 2, 3, 4, 6, 8, 9, 77, -1, 2**40])
                hdr = struct.pack("<4q", op, rnd.choice([0, 1, -5, 2**62, 10**9]), rnd.randrange(-3, 100), 0)
                s.sendall(hdr[:rnd.choice([32, 32, 32, 7])])
                s.sendall(bytes(rnd.randrange(256) for _ in range(rnd.choice([0, 0, 13, 256, 512]))))
                try:
                    s.recv(4096)
                except (socket.timeout, ConnectionError):
                    pass
        except OSError:
            pass
        s.close()
    c = ps.channel(0)
    w0 = torch.arange(64, dtype=torch.float32)
    c.call(0, OP_INIT, 5, payload=torch.cat([w0, w0]))
    hdr, w = c.call(0, OP_PULL, reply_numel=64)
    assert int(hdr[1]) == 5 and torch.equal(w, w0)
    c.call(0, OP_DONE)
    ps.close()
    assert ps.rc == 0


def test_one_of_two_aggregation_drops_the_slower_replica():
    """replicas_to_aggregate=1, total_num_replicas=2 (backup-worker mode): the first fresh gradient moves the step,
    the slower replica's gradient for the same step is stale; two tokens per step keep both workers going."""
    ps = _Ps(numel=2, n_workers=2, n_agg=1, use_ema=0)
    a, b = ps.channel(0), ps.channel(1)
    a.call(0, OP_INIT, 0, payload=torch.zeros(4))
    a.call(0, OP_SYNC_INIT, 1)
    g = torch.tensor([1.0, 2.0])
    hdr, _ = a.call(0, OP_PUSH_SYNC, _lr_bits(1.0), 0, payload=g)
    assert int(hdr[1]) == 1
    hdr, _ = b.call(0, OP_PUSH_SYNC, _lr_bits(1.0), 0, payload=10 * g)
    assert int(hdr[1]) == 1 and ps.stats()["dropped"] == 1
    _, w = a.call(0, OP_PULL, reply_numel=2)
    assert torch.equal(w, -g)
    toks = sorted(int(c.call(0, OP_TOKEN)[0][1]) for c in (a, b, a))
    assert toks == [0, 1, 1]  # one pre-loaded token + max(total, aggregate) = 2 for the new step
    a.call(0, OP_DONE), b.call(0, OP_DONE)
    ps.close()
