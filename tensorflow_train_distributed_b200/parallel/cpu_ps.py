"""CPU parameter server over gloo — BASELINE config 1 (1 ps + 2 workers on localhost, no GPU).

This is the tier where the reference's process model is kept literally: ``--job_name=ps`` processes hold the
variables ``replica_device_setter`` would give them (round-robin in creation order, ``global_step`` first —
distribute_training.py:116, :186-190) and serve pulls and pushes; workers compute gradients with a plain PyTorch
CPU model and talk to every ps each step:

async (:150-152)   pull -> fwd/bwd -> push(grads, lr) ; the ps applies at once, ps0 bumps ``global_step``
sync  (:142-148)   ``SyncReplicasOptimizer`` semantics: per-ps gradient accumulator that drops
                   gradients tagged with a stale ``local_step``, mean over ``replicas_to_aggregate`` fresh
                   gradients, ONE ``global_step`` increment per aggregate, a token queue on ps0 that hands every
                   worker the new global step (pre-loaded with ``replicas_to_aggregate`` tokens, so each
                   worker's second gradient is stale and dropped — local steps run one ahead).

Transport: ``torch.distributed`` gloo send/recv; a ps is a single-threaded event loop on ``recv(any source)``.
Message = int64 header ``[op, a, b, c]`` (tag 0) followed by an optional float32 payload (tag 1).
"""
from __future__ import annotations

import collections
import datetime as _dt
import time
from dataclasses import dataclass
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.distributed as dist

from ..cluster import ClusterSpec, TaskInfo
from ..config import TrainConfig
from ..session import TransportError
from ..utils.schedule import ema_decay, exponential_decay
from . import sharding
from .sharding import EMA_SUFFIX, GLOBAL_STEP, ParamLayout, VariableSpec

OP_READY, OP_INIT, OP_PULL, OP_PUSH_ASYNC, OP_PUSH_SYNC, OP_TOKEN, OP_GET_STATE, OP_DONE, OP_SYNC_INIT = range(9)
TAG_HDR, TAG_PAYLOAD = 0, 1


def init_gloo(task: TaskInfo, timeout_s: float = 120.0) -> None:
    if dist.is_initialized():
        return
    dist.init_process_group("gloo", init_method=f"tcp://{task.master_addr}:{task.master_port}", rank=task.rank,
                            world_size=task.world_size, timeout=_dt.timedelta(seconds=timeout_s))


def _send(dst: int, op: int, a: int = 0, b: int = 0, c: int = 0, payload: Optional[torch.Tensor] = None) -> None:
    dist.send(torch.tensor([op, a, b, c], dtype=torch.int64), dst=dst, tag=TAG_HDR)
    if payload is not None and payload.numel():
        dist.send(payload.contiguous(), dst=dst, tag=TAG_PAYLOAD)


def _recv_payload(src: int, n: int) -> torch.Tensor:
    t = torch.empty(n, dtype=torch.float32)
    if n:
        dist.recv(t, src=src, tag=TAG_PAYLOAD)
    return t


# =====================================================================================================
# model on the worker (plain PyTorch CPU; TF-layout kernels [in, out])
# =====================================================================================================
class CpuMlp:
    """forward_propagation + loss + in_top_k of the reference (:39-63, :77-87, :96) with autograd."""

    def __init__(self, layout: ParamLayout, activation: str, dropout_rate: float, seed: int):
        self.layout, self.activation, self.dropout_rate = layout, activation, dropout_rate
        self.gen = torch.Generator().manual_seed(seed)

    def logits(self, leaves: Dict[str, torch.Tensor], x: torch.Tensor, training: bool = True) -> torch.Tensor:
        """``forward_propagation`` (:39-63); the reference hard-codes training=True (:94), evaluation turns dropout off."""
        h = x
        groups = self.layout.groups
        for i, g in enumerate(groups):
            k = leaves[g.kernel_name]
            if k.dim() == 4:  # LeNet conv stage (config 5): HWIO kernel, SAME padding for the first, 2x2 max-pool
                kh, kw, cin, cout = k.shape
                if h.dim() == 2:
                    side = int(round((h.shape[1] // cin) ** 0.5))
                    h = h.view(-1, side, side, cin).permute(0, 3, 1, 2)  # NHWC batch -> NCHW
                h = torch.nn.functional.conv2d(h, k.permute(3, 2, 0, 1), leaves[g.bias_name],
                                               padding=(kh // 2) if i == 0 else 0)
                h = torch.nn.functional.elu(h) if self.activation == "elu" else (
                    torch.relu(h) if self.activation == "relu" else h)
                h = torch.nn.functional.max_pool2d(h, 2)
                continue
            if h.dim() == 4:
                h = h.permute(0, 2, 3, 1).reshape(h.shape[0], -1)  # TF flatten order [H, W, C]
            h = h @ k + leaves[g.bias_name]
            if i < len(groups) - 1:
                if self.activation == "elu":
                    h = torch.nn.functional.elu(h)
                elif self.activation == "relu":
                    h = torch.relu(h)
                if self.dropout_rate > 0 and training:
                    keep = 1.0 - self.dropout_rate
                    mask = (torch.rand(h.shape, generator=self.gen) < keep).to(h.dtype) / keep
                    h = h * mask
        return h

    def loss_and_grads(self, params: Dict[str, torch.Tensor], x: torch.Tensor, y: torch.Tensor):
        leaves = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
        h = self.logits(leaves, x, training=True)  # always training=True (:94)
        loss = torch.nn.functional.cross_entropy(h, y)
        loss.backward()
        zl = h.gather(1, y.view(-1, 1))
        acc = ((h > zl).sum(1) < 1).float().mean()  # in_top_k(logits, labels, 1)
        return float(loss.detach()), float(acc), {k: v.grad for k, v in leaves.items()}


def he_init(layout: ParamLayout, seed: int) -> Dict[str, np.ndarray]:
    """variance_scaling_initializer(): truncated normal, sigma = sqrt(1.3 * 2 / fan_in); zero biases (:49, A.5)."""
    g = torch.Generator().manual_seed(seed)
    out: Dict[str, np.ndarray] = {}
    for grp in layout.groups:
        sigma = (2.6 / grp.rows) ** 0.5
        w = torch.empty(grp.rows, grp.cols)
        torch.nn.init.trunc_normal_(w, mean=0.0, std=sigma, a=-2 * sigma, b=2 * sigma, generator=g)
        out[grp.kernel_name] = w.reshape(grp.kernel_shape).numpy()
        out[grp.bias_name] = np.zeros(grp.cols, dtype=np.float32)
    return out


# =====================================================================================================
# placement helpers
# =====================================================================================================
@dataclass
class Placement:
    variables: List[VariableSpec]
    owner: Dict[str, int]

    def names_on(self, ps: int) -> List[str]:
        return [v.name for v in self.variables if v.trainable and self.owner[v.name] == ps]

    def numel_on(self, ps: int) -> int:
        return sum(v.numel for v in self.variables if v.trainable and self.owner[v.name] == ps)


def make_placement(layout: ParamLayout, n_ps: int) -> Placement:
    variables = layout.variables(with_global_step=True)
    return Placement(variables, sharding.round_robin(variables, n_ps))


def _flatten(names: List[str], tensors: Dict[str, torch.Tensor]) -> torch.Tensor:
    if not names:
        return torch.empty(0)
    return torch.cat([tensors[n].reshape(-1).float() for n in names])


def _unflatten(names: List[str], specs: Dict[str, VariableSpec], flat: torch.Tensor) -> Dict[str, torch.Tensor]:
    out, off = {}, 0
    for n in names:
        k = specs[n].numel
        out[n] = flat[off:off + k].view(specs[n].shape).clone()
        off += k
    return out


# =====================================================================================================
# the ps process
# =====================================================================================================
class PsServer:
    """``server.join()`` of the reference (:179-181) — but it returns once every worker has said goodbye."""

    def __init__(self, cfg: TrainConfig, cluster: ClusterSpec, task: TaskInfo, layout: ParamLayout):
        self.cfg, self.cluster, self.task, self.layout = cfg, cluster, task, layout
        self.ps_index = task.index
        self.n_workers = cluster.num_tasks("worker")
        self.place = make_placement(layout, cluster.num_tasks("ps"))
        self.specs = {v.name: v for v in self.place.variables}
        self.names = self.place.names_on(self.ps_index)
        self.numel = self.place.numel_on(self.ps_index)
        self.owns_global_step = self.place.owner[GLOBAL_STEP] == self.ps_index
        self.w = torch.zeros(self.numel)
        self.ema = torch.zeros(self.numel)
        self.reg_mask = torch.zeros(self.numel)
        off = 0
        for n in self.names:
            k = self.specs[n].numel
            if self.specs[n].regularize:
                self.reg_mask[off:off + k] = 1.0
            off += k
        self.global_step = 0
        self.initialized = False
        # sync state
        self.n_agg = cfg.aggregate_n
        self.acc = torch.zeros(self.numel)
        self.acc_count = 0
        self.acc_step = 0
        self.tokens: collections.deque = collections.deque()
        self.pending_tokens: collections.deque = collections.deque()
        self.done_workers = set()
        self.stats = {"applied": 0, "dropped_stale": 0, "pulls": 0}

    # -- update rule: what TF places on the ps (ApplyGradientDescent, :150-152) + L1/L2 + EMA ---------
    def _apply(self, grad: torch.Tensor, lr: float, gs_after: int) -> None:
        cfg = self.cfg
        if cfg.regularizer == "l2":
            grad = grad + cfg.regularizer_scale * self.w * self.reg_mask
        elif cfg.regularizer == "l1":
            grad = grad + cfg.regularizer_scale * torch.sign(self.w) * self.reg_mask
        self.w -= lr * grad
        if cfg.use_moving_average:
            d = ema_decay(cfg.moving_average_decay, gs_after)
            self.ema -= (1.0 - d) * (self.ema - self.w)
        self.stats["applied"] += 1

    def _issue_tokens(self) -> None:
        while self.pending_tokens and self.tokens:
            dst = self.pending_tokens.popleft()
            _send(dst, OP_TOKEN, self.tokens.popleft(), self.global_step)

    def serve(self) -> Dict[str, int]:
        hdr = torch.zeros(4, dtype=torch.int64)
        self._last_msg = time.time()
        idle_limit = float(self.cfg.ps_idle_timeout_secs)
        if idle_limit > 0:
            # watchdog thread: a crashed worker never says goodbye, and a blocking recv(any source) cannot be
            # interrupted — so a side thread ends the process instead of serving nobody forever
            import os
            import threading

            def _watch():
                while True:
                    time.sleep(min(0.5, idle_limit / 4))
                    if self.initialized and time.time() - self._last_msg > idle_limit:
                        print(f"ps {self.ps_index} done: {dict(self.stats, idle_timeout=1, global_step=self.global_step)}",
                              flush=True)
                        os._exit(0)
            threading.Thread(target=_watch, daemon=True).start()
        while len(self.done_workers) < self.n_workers:
            src = dist.recv(hdr, src=None, tag=TAG_HDR)
            self._last_msg = time.time()
            op, a, b, c = (int(v) for v in hdr)
            if op == OP_READY:
                _send(src, OP_READY, 1 if self.initialized else 0, self.global_step)
            elif op == OP_INIT:  # chief: a = global_step; payload = [w | ema]
                flat = _recv_payload(src, 2 * self.numel)
                self.w, self.ema = flat[:self.numel].clone(), flat[self.numel:].clone()
                self.global_step, self.acc_step = a, a
                self.acc.zero_()
                self.acc_count = 0
                self.initialized = True
                _send(src, OP_INIT, 1)
            elif op == OP_SYNC_INIT:  # chief sync hook: pre-load a tokens carrying the current global step
                if self.owns_global_step:
                    for _ in range(a):
                        self.tokens.append(self.global_step)
                    self._issue_tokens()
                _send(src, OP_SYNC_INIT, 1)
            elif op == OP_PULL:
                self.stats["pulls"] += 1
                _send(src, OP_PULL, self.global_step, payload=self.w)
            elif op == OP_PUSH_ASYNC:  # a = lr bits, payload = grads; applied immediately, unlocked (:150)
                grad = _recv_payload(src, self.numel)
                lr = float(np.int64(a).view(np.float64))
                if self.owns_global_step:
                    self.global_step += 1
                self._apply(grad, lr, b + 1)
                _send(src, OP_PUSH_ASYNC, self.global_step)
            elif op == OP_PUSH_SYNC:  # a = lr bits, b = local_step tag
                grad = _recv_payload(src, self.numel)
                lr = float(np.int64(a).view(np.float64))
                if b < self.acc_step:
                    self.stats["dropped_stale"] += 1  # ConditionalAccumulator drops stale gradients silently
                else:
                    self.acc += grad
                    self.acc_count += 1
                    if self.acc_count >= self.n_agg:  # take_grad(n): mean of n fresh gradients
                        new_gs = self.acc_step + 1
                        self._apply(self.acc / float(self.acc_count), lr, new_gs)
                        self.acc.zero_()
                        self.acc_count = 0
                        self.acc_step = new_gs
                        if self.owns_global_step:
                            self.global_step = new_gs
                            n_tok = max(self.cfg.total_num_replicas or self.n_workers, self.n_agg)
                            for _ in range(n_tok):
                                self.tokens.append(new_gs)
                            self._issue_tokens()
                _send(src, OP_PUSH_SYNC, self.global_step)
            elif op == OP_TOKEN:
                if self.done_workers and not self.tokens:
                    # a worker has already left, so no further aggregate can complete: answer at once
                    _send(src, OP_TOKEN, self.global_step, self.global_step)
                else:
                    self.pending_tokens.append(src)
                    self._issue_tokens()
            elif op == OP_GET_STATE:
                _send(src, OP_GET_STATE, self.global_step, payload=torch.cat([self.w, self.ema]))
            elif op == OP_DONE:
                self.done_workers.add(src)
                # nobody will complete another aggregate: release blocked token requests so they can stop
                while self.pending_tokens:
                    _send(self.pending_tokens.popleft(), OP_TOKEN, self.global_step, self.global_step)
                _send(src, OP_DONE, 1)
            else:
                raise RuntimeError(f"ps{self.ps_index}: unknown op {op} from rank {src}")
        return dict(self.stats, global_step=self.global_step)


# =====================================================================================================
# the worker-side trainer
# =====================================================================================================
@dataclass
class CpuStepValues:
    loss: float
    accuracy: float
    global_step: int


class PsWorkerTrainer:
    """Worker half of the CPU parameter server; implements the trainer interface of the session."""

    every_rank_restores = False

    def __init__(self, cfg: TrainConfig, cluster: ClusterSpec, task: TaskInfo, layout: ParamLayout, channel=None):
        self.cfg, self.cluster, self.task, self.layout = cfg, cluster, task, layout
        self.chan = channel  # None: gloo send/recv; else an object with call(ps, op, a, b, c, payload, reply_numel)
        self.n_ps = cluster.num_tasks("ps")
        self.place = make_placement(layout, self.n_ps)
        self.specs = {v.name: v for v in self.place.variables}
        self.ps_names = [self.place.names_on(k) for k in range(self.n_ps)]
        self.ps_numel = [self.place.numel_on(k) for k in range(self.n_ps)]
        self.gs_ps = self.place.owner[GLOBAL_STEP]
        # push order: the ps holding global_step goes LAST, so when it completes an aggregate every other ps
        # has already accumulated all N gradients (strict version of TF's control dependencies)
        self.push_order = [k for k in range(self.n_ps) if k != self.gs_ps] + [self.gs_ps]
        self.model = CpuMlp(layout, cfg.activation, cfg.dropout_rate, cfg.seed + 7919 * task.index)
        self.local_step = 0      # sync_rep_local_step
        self._gs = 0
        self._closed = False

    # -- rpc helpers ---------------------------------------------------------------------------------------
    def _call(self, ps: int, op: int, a: int = 0, b: int = 0, c: int = 0, payload=None, reply_numel: int = 0):
        if self.chan is not None:
            return self.chan.call(ps, op, a, b, c, payload, reply_numel)
        try:
            _send(ps, op, a, b, c, payload)
            hdr = torch.zeros(4, dtype=torch.int64)
            dist.recv(hdr, src=ps, tag=TAG_HDR)
            body = _recv_payload(ps, reply_numel) if reply_numel else None
        except RuntimeError as exc:  # gloo surfaces a dead peer as RuntimeError
            raise TransportError(f"lost parameter server {ps}: {exc}") from exc
        return hdr, body

    # -- session interface ----------------------------------------------------------------------------------
    def initialize(self, restore: Optional[Dict[str, np.ndarray]] = None) -> None:
        cfg = self.cfg
        if self.task.is_chief:  # chief: init_op or Saver.restore, executed "on the ps" 
            values = he_init(self.layout, cfg.seed)
            gs = 0
            ema = dict(values)
            if restore is not None:
                for n in values:
                    values[n] = np.asarray(restore[n], dtype=np.float32)
                    ema[n] = np.asarray(restore.get(n + EMA_SUFFIX, restore[n]), dtype=np.float32)
                gs = int(np.asarray(restore.get(GLOBAL_STEP, 0)))
            tv = {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in values.items()}
            te = {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in ema.items()}
            for k in range(self.n_ps):
                flat = torch.cat([_flatten(self.ps_names[k], tv), _flatten(self.ps_names[k], te)])
                self._call(k, OP_INIT, gs, payload=flat)
            if cfg.sync_replicas:  # make_session_run_hook(is_chief): pre-load the token queue (A.2)
                self._call(self.gs_ps, OP_SYNC_INIT, cfg.aggregate_n)
            self._gs = gs
        else:  # WorkerSessionCreator.wait_for_session: poll until the chief has initialised
            deadline = time.time() + cfg.rendezvous_timeout_secs
            while True:
                ready = True
                for k in range(self.n_ps):
                    hdr, _ = self._call(k, OP_READY)
                    ready = ready and int(hdr[1]) == 1
                    if k == self.gs_ps:
                        self._gs = int(hdr[2])
                if ready:
                    break
                if time.time() > deadline:
                    raise TransportError("timed out waiting for the chief to initialise the variables")
                time.sleep(0.05)
        self.local_step = self._gs  # SyncReplicasOptimizerHook: local_step <- global_step

    def reconnect(self, restore=None) -> None:
        self.initialize(restore)

    def _pull(self) -> Dict[str, torch.Tensor]:
        params: Dict[str, torch.Tensor] = {}
        for k in range(self.n_ps):
            hdr, flat = self._call(k, OP_PULL, reply_numel=self.ps_numel[k])
            if k == self.gs_ps:
                self._gs = int(hdr[1])
            params.update(_unflatten(self.ps_names[k], self.specs, flat))
        return params

    def train_step(self, x, y) -> CpuStepValues:
        cfg = self.cfg
        x = torch.as_tensor(np.asarray(x), dtype=torch.float32)
        y = torch.as_tensor(np.asarray(y), dtype=torch.int64)
        params = self._pull()  # M1: parameter pull + global_step for the lr
        gs_seen = self._gs
        loss, acc, grads = self.model.loss_and_grads(params, x, y)
        lr = exponential_decay(cfg.initial_learning_rate, gs_seen, cfg.decay_steps, cfg.learning_rate_decay_factor,
                               cfg.staircase)
        lr_bits = int(np.float64(lr).view(np.int64))
        op = OP_PUSH_SYNC if cfg.sync_replicas else OP_PUSH_ASYNC
        tag = self.local_step if cfg.sync_replicas else gs_seen
        for k in self.push_order:  # M2 / M3: gradient push
            hdr, _ = self._call(k, op, lr_bits, tag, payload=_flatten(self.ps_names[k], grads))
            if k == self.gs_ps:
                self._gs = int(hdr[1])
        if cfg.sync_replicas:  # M4: block on the token queue, adopt the token as the new local step
            hdr, _ = self._call(self.gs_ps, OP_TOKEN)
            self.local_step = int(hdr[1])
            self._gs = max(self._gs, int(hdr[2]))
        else:
            self.local_step += 1
        return CpuStepValues(loss, acc, self._gs)

    def global_step(self) -> int:
        return self._gs

    def export_variables(self) -> Dict[str, np.ndarray]:
        out: Dict[str, np.ndarray] = {}
        for k in range(self.n_ps):
            hdr, flat = self._call(k, OP_GET_STATE, reply_numel=2 * self.ps_numel[k])
            n = self.ps_numel[k]
            for name, t in _unflatten(self.ps_names[k], self.specs, flat[:n]).items():
                out[name] = t.numpy()
            if self.cfg.use_moving_average:
                for name, t in _unflatten(self.ps_names[k], self.specs, flat[n:]).items():
                    out[name + EMA_SUFFIX] = t.numpy()
            if k == self.gs_ps:
                self._gs = int(hdr[1])
        out[GLOBAL_STEP] = np.asarray(self._gs, dtype=np.int64)
        return out

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        for k in range(self.n_ps):
            try:
                self._call(k, OP_DONE)
            except TransportError:
                pass
        if self.chan is not None:
            self.chan.close()
