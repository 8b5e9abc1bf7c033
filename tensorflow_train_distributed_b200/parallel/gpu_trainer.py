"""GPU worker: owns a parameter-store shard, a model and the captured training step.

In the reference a worker's step is ``mon_sess.run([train_op, loss, accuracy, global_step], feed_dict)``
(distribute_training.py:224-226): ship the batch to the runtime, pull parameters from the ps over gRPC,
compute, push gradients back, fetch three scalars. Here the same step is one CUDA-graph replay on the GPU that
also owns 1/N of the parameters; the host copies the batch in, and reads the 64-byte state block back — per
step on the end-to-end path, or only at log cadence (:229) when the caller does not need every value.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from ..config import TrainConfig
from ..models import mlp as mlp_model
from ..ops import _native as N
from .param_store import GpuParamStore, Hyper
from ..utils.tracing import nvtx_range
from .sharding import ParamLayout


@dataclass
class StepValues:
    """What the reference's ``mon_sess.run`` returns (:225): loss, accuracy, global_step."""
    loss: float
    accuracy: float
    global_step: int
    local_step: int
    learning_rate: float
    pending: int = 0  # --fetch_lag: steps enqueued after the one these values belong to


def hyper_from_config(cfg: TrainConfig, n_replicas: int) -> Hyper:
    return Hyper(lr0=cfg.initial_learning_rate, decay_rate=cfg.learning_rate_decay_factor,
                 decay_steps=cfg.decay_steps, staircase=cfg.staircase, regularizer=cfg.regularizer,
                 reg_scale=cfg.regularizer_scale,
                 ema_decay=cfg.moving_average_decay if cfg.use_moving_average else 0.0, ema_warmup=True,
                 grad_scale=1.0 / n_replicas)


class GpuTrainer:
    """One process per GPU. ``comm`` (a SymmetricComm) joins several of them into one parameter server."""

    RING = 64  # pinned result slots for asynchronous per-step fetches
    # Every worker reads the checkpoint itself (reference :208-215: after the chief's restore EVERY variable on
    # EVERY ps holds the restored value). Each GPU is the ps of one flat range and keeps its own global_step, so
    # a restore that reached only the chief would leave the other owners on their seed-initialised shards.
    every_rank_restores = True

    def __init__(self, cfg: TrainConfig, device_index: int = 0, comm=None):
        if not torch.cuda.is_available():
            raise RuntimeError("GpuTrainer needs a CUDA device; use --device cpu (tcp or gloo transport) on CPU-only machines")
        self.cfg, self.comm = cfg, comm
        self.device = torch.device("cuda", device_index)
        torch.cuda.set_device(self.device)
        self.world = comm.world_size if comm is not None else 1
        self.rank = comm.rank if comm is not None else 0
        self.sync = cfg.sync_replicas or self.world == 1
        self.stream = torch.cuda.Stream(device=self.device)
        self.layout: ParamLayout = self._build_layout()
        if self.sync and cfg.replicas_to_aggregate not in (0, self.world):
            raise ValueError(f"--replicas_to_aggregate={cfg.replicas_to_aggregate}: the GPU sync path aggregates exactly one "
                             f"gradient per replica ({self.world}); N-of-M aggregation exists in the CPU tier (--device cpu)")
        n_agg = (cfg.replicas_to_aggregate or self.world) if self.sync else 1
        self.hyper = hyper_from_config(cfg, n_agg)
        with torch.cuda.stream(self.stream):
            lenet = cfg.model == "lenet"
            cluster_tail = cfg.fused_tail and cfg.tail_impl == "cluster"
            if lenet:  # dense head = groups 2.. (fc1, fc2, output); its first layer stays on the GEMM path
                tail = mlp_model.tail_groups([120, 84, cfg.num_classes], cfg.batch_size, 2, cluster_tail)
            else:
                tail = mlp_model.tail_groups(cfg.layer_sizes, cfg.batch_size, 0, cluster_tail)
            self.store = GpuParamStore(self.layout, self.hyper, self.device, comm=comm, need_w_rm_first=lenet,
                                       pull_fused_first=self.sync and self.world > 1 and cfg.pull_fused and not lenet,
                                       tail_groups=tail)
            self.model = self._build_model()
            if cfg.step_impl == "column" and not getattr(self.model, "use_column", False):
                raise ValueError("--step_impl=column: this model / configuration does not fit the column kernel (narrow MLP "
                                 "stacks only: hidden widths <= 256, <= 32 classes, no --fused_tail / --pull_fused)")
            # ring of step states written by the device itself (TtdHyper.host_mirror): slot = step index % RING
            self._pinned_state = torch.zeros(self.RING, 8, dtype=torch.int64).pin_memory()
            self._pinned_probe = torch.zeros(8, dtype=torch.int64).pin_memory()  # explicit read-backs (slot -1)
            self.store.set_host_mirror(self._pinned_state.data_ptr(), self.RING)
            # Host-fed batches land in a two-slot device pool; the captured step picks slot local_step % 2 on the
            # device, so the H2D copy of batch s+1 (copy stream) overlaps the compute of batch s (one graph).
            B, F = self.model.x.shape
            self._feed_x = torch.zeros(2, B, F, device=self.device, dtype=torch.float32)
            self._feed_y = torch.zeros(2, B, device=self.device, dtype=torch.int64)
            self._feed_prep = self._attach_feed_pool()
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.lib = N.load()
        self.lib.ttd_set_barrier_timeout_ms(int(max(1.0, cfg.barrier_timeout_secs) * 1000))
        self._feeder = self.lib.ttd_feeder_create(
            self.stream.cuda_stream, self.copy_stream.cuda_stream, self._feed_x[0].data_ptr(), self._feed_x[1].data_ptr(),
            self._feed_y[0].data_ptr(), self._feed_y[1].data_ptr(), self._feed_x[0].numel() * 4, self._feed_y[0].numel() * 8)
        if not self._feeder:
            raise N.NativeError("ttd_feeder_create failed: " + self.lib.ttd_last_error().decode(errors="replace"))
        self._configure_feeder_prep()
        self._host_keep = [None, None]
        import collections as _collections
        self._inflight = _collections.deque()
        self._last_values: Optional[StepValues] = None
        self._pinned_np = self._pinned_state.numpy()  # host views of the pinned buffers (polled, never copied)
        self._pinned_u32 = self._pinned_np.view(np.uint32)
        self._pinned_np_probe = self._pinned_probe.numpy()
        self._expect = [0] * self.RING
        self._graph_exec: Optional[int] = None
        self._user_pool = False
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._launches_per_step = 0
        self._slot = 0
        self._steps_enqueued = 0
        self._initialized = False
        self._closed = False

    # -------------------------------------------------------------------------------------------
    def _build_layout(self) -> ParamLayout:
        cfg = self.cfg
        if cfg.model in ("mlp", "wide_mlp"):
            return mlp_model.build_layout(cfg.input_dim, cfg.layer_sizes)
        if cfg.model == "lenet":
            from ..models import lenet
            return lenet.build_layout(cfg.num_classes)
        raise ValueError(cfg.model)

    def _build_model(self):
        cfg = self.cfg
        if cfg.model in ("mlp", "wide_mlp"):
            return mlp_model.GpuMlp(self.store, cfg.batch_size, activation=cfg.activation,
                                    dropout_rate=cfg.dropout_rate, seed=cfg.seed + 7919 * self.rank, pdl=cfg.pdl,
                                    chain=cfg.fused_tail and cfg.tail_impl == "chain",
                                    net=cfg.fused_tail and cfg.tail_impl == "net",
                                    column=cfg.step_impl != "layers" and not cfg.fused_tail)
        from ..models import lenet
        return lenet.GpuLeNet(self.store, cfg.batch_size, activation=cfg.activation,
                              dropout_rate=cfg.dropout_rate, seed=cfg.seed + 7919 * self.rank, pdl=cfg.pdl,
                              chain=cfg.fused_tail and cfg.tail_impl == "chain",
                              net=cfg.fused_tail and cfg.tail_impl == "net")

    # -------------------------------------------------------------------------------------------
    def initialize(self, restore: Optional[Dict[str, np.ndarray]] = None) -> None:
        """Chief-init or restore (MonitoredTrainingSession bring-up, :209-215), then capture the step graph."""
        with torch.cuda.stream(self.stream):
            self.store.init(self.cfg.seed)
            # one eager step warms every kernel (function attributes, tensor maps, module load) ...
            self._launches_per_step = self.model.launch_step(sync=self.sync)
            self.stream.synchronize()
            if self.comm is not None:
                self.comm.barrier()
            # ... then the real initial state is laid down again
            self.store.state.zero_()
            if self.comm is None:
                self.store.grads.zero_()
            self.store.init(self.cfg.seed)
            if restore is not None:
                self.store.import_variables(restore)
            self.stream.synchronize()
            if self.comm is not None:
                self.comm.reset_flags(self.store)
                self.comm.barrier()
                self._check_replicas_agree(restore is not None)
        self._steps_enqueued = 0  # mirrors the device's local_step (which slot of the feed pool a step reads)
        self._pinned_state.zero_()  # entries of the warm-up step (or of an earlier run) must not pass for new ones
        if self.cfg.cuda_graph:
            self._capture()
        self._inflight.clear()
        self._last_values = self.read_values()  # global_step of a restored checkpoint, zeros otherwise
        self._initialized = True

    def _check_replicas_agree(self, restored: bool) -> None:
        """All replicas must start from the same global_step and the same parameters (a replica that missed the
        restore would otherwise train on with a different learning rate and stop at a different step)."""
        import torch.distributed as dist
        digest = torch.stack([self.store.state[0].double(), torch.tensor(1.0 if restored else 0.0, device=self.device,
                                                                         dtype=torch.float64),
                              self.store.w.double().sum(), self.store.ema.double().sum()])
        all_d = [torch.zeros_like(digest) for _ in range(self.world)]
        dist.all_gather(all_d, digest)
        if not all(torch.equal(all_d[0], d) for d in all_d):
            raise RuntimeError("replicas disagree after initialisation (global_step, restored?, sum(w), sum(ema)): "
                               + "; ".join(str(d.tolist()) for d in all_d))

    def reconnect(self, restore: Optional[Dict[str, np.ndarray]] = None) -> None:
        """Session recovery (``_RecoverableSession``): lay the state down again — from the newest checkpoint when
        there is one — and re-capture the step. All replicas have to call it together (it contains barriers)."""
        self.stream.synchronize()
        self._graph = None
        self._graph_exec = None
        self.initialize(restore=restore)

    def _capture(self) -> None:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=self.stream):
            self.model.launch_step(sync=self.sync)
        self._graph = g
        self._graph_exec = int(g.raw_cuda_graph_exec())  # launched by the native feeder (ttd_feeder_step)

    def _attach_feed_pool(self):
        """Point the model at the two-slot feed pool. Models that want their input pre-converted (column mode of
        the MLP) return the conversion targets; the native feeder then converts each slot on the copy stream."""
        set_feed = getattr(self.model, "set_feed_pool", None)
        if set_feed is not None:
            return set_feed(self._feed_x, self._feed_y)
        self.model.set_device_dataset(self._feed_x, self._feed_y)
        return None

    def _configure_feeder_prep(self) -> None:
        prep = self._feed_prep
        B, F = self.model.x.shape
        if prep is None:
            N.check(self.lib.ttd_feeder_set_prep(self._feeder, B, F, None, 0, None, 0), "ttd_feeder_set_prep")
        else:
            xb, ldx, xT, ldt = prep
            N.check(self.lib.ttd_feeder_set_prep(self._feeder, B, F, xb.data_ptr(), ldx, xT.data_ptr(), ldt),
                    "ttd_feeder_set_prep")

    def use_device_dataset(self, x_pool, labels_pool) -> None:
        """Switch between a device-resident data set ([P,B,F] / [P,B], batch picked on the device from the
        local step) and host-fed batches (``None``); the step graph is re-captured for the new source."""
        self.stream.synchronize()
        self.copy_stream.synchronize()
        self._user_pool = x_pool is not None
        with torch.cuda.stream(self.stream):
            if x_pool is None:
                self._feed_prep = self._attach_feed_pool()
                self._configure_feeder_prep()
            else:
                self.model.set_device_dataset(x_pool, labels_pool)
            self.stream.synchronize()
        if self._graph is not None:
            self._capture()

    # -------------------------------------------------------------------------------------------
    @property
    def launches_per_step(self) -> int:
        return self._launches_per_step

    def _host_batch(self, x, y):
        x = torch.as_tensor(x)
        y = torch.as_tensor(y)
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.to(torch.float32).contiguous()
        if y.dtype != torch.int64 or not y.is_contiguous():
            y = y.to(torch.int64).contiguous()
        if x.numel() != self._feed_x[0].numel() or y.numel() != self._feed_y[0].numel():
            raise ValueError(f"batch of shape {tuple(x.shape)} / {tuple(y.shape)} does not match the model's "
                             f"{tuple(self._feed_x[0].shape)} / {tuple(self._feed_y[0].shape)}")
        return x, y

    def load_batch(self, x, y) -> Tuple[int, int]:
        """Host -> device copy of the batch for the NEXT step, on the copy stream, into the feed-pool slot that
        step will read; returns (bytes_x, bytes_y). Pinned sources make the copy asynchronous."""
        x, y = self._host_batch(x, y)
        slot = self._steps_enqueued % 2
        N.check(self.lib.ttd_feeder_feed(self._feeder, x.data_ptr(), y.data_ptr(), self._steps_enqueued), "ttd_feeder_feed")
        self._host_keep[slot] = (x, y)  # the copy may still be reading them
        return x.numel() * x.element_size(), y.numel() * y.element_size()

    def wait_batch(self) -> Optional[int]:
        """Make the compute stream wait for the batch fed by ``load_batch``; returns its feed-pool slot."""
        slot = self.lib.ttd_feeder_before_step(self._feeder)
        return slot if slot >= 0 else None

    def run_step(self) -> None:
        """Enqueue one training step (on the batch fed for it, or on the device-resident data set)."""
        if self._graph_exec is not None:
            N.check(self.lib.ttd_feeder_step(self._feeder, self._graph_exec, None, None, 0, None, None), "ttd_feeder_step")
        else:
            self.wait_batch()
            with torch.cuda.stream(self.stream):
                self.model.launch_step(sync=self.sync)
            N.check(self.lib.ttd_feeder_after_step(self._feeder), "ttd_feeder_after_step")
        self._steps_enqueued += 1

    def step_async(self, x, y, fetch: bool = True) -> Optional[int]:
        """Feed one host batch, run the step and (optionally) start the read-back of its 64-byte state — one
        native call when the step is a captured graph. Returns the pinned ring slot for ``read_slot``."""
        if self._graph_exec is None:
            self.load_batch(x, y)
            self.run_step()
            if not fetch:
                return None
            with torch.cuda.stream(self.stream):
                return self.fetch_async()
        x, y = self._host_batch(x, y)
        self._host_keep[self._steps_enqueued % 2] = (x, y)
        slot = self._steps_enqueued % self.RING  # where the update kernel mirrors this step's state
        N.check(self.lib.ttd_feeder_step(self._feeder, self._graph_exec, x.data_ptr(), y.data_ptr(), self._steps_enqueued,
                                         None, None), "ttd_feeder_step")
        self._steps_enqueued += 1
        self._expect[slot] = self._steps_enqueued
        return slot if fetch else None

    def fetch_async(self) -> int:
        """Enqueue an explicit device -> host copy of the current step state (slot -1 of ``read_slot``)."""
        self._pinned_probe.copy_(self.store.state, non_blocking=True)
        return -1

    def read_slot(self, slot: int) -> StepValues:
        buf = self._pinned_probe if slot < 0 else self._pinned_state[slot]
        st = N.StepState.from_buffer_copy(buf.numpy().tobytes())
        if slot >= 0 and st.local_step != self._expect[slot]:
            # the device's step counter was advanced behind the trainer's back (a direct store.apply_update):
            # the mirror slot is not the one predicted, read the state explicitly instead
            with torch.cuda.stream(self.stream):
                self.fetch_async()
            self.stream.synchronize()
            st = N.StepState.from_buffer_copy(self._pinned_probe.numpy().tobytes())
        return StepValues(st.metrics_out[0], st.metrics_out[1], st.global_step, st.local_step, st.lr_last)

    def train_step(self, x, y, fetch: bool = True) -> Optional[StepValues]:
        """The user-facing step: feed one batch, run, and (optionally) fetch loss / accuracy / global_step."""
        with nvtx_range("ttd.step.enqueue"):
            slot = self.step_async(x, y, fetch)
        if slot is None:
            return None
        lag = self.cfg.fetch_lag if (self._graph_exec is not None and slot >= 0) else 0
        if lag <= 0:
            with nvtx_range("ttd.step.fetch"):
                self.stream.synchronize()
                return self.read_slot(slot)
        # pipelined fetch (--fetch_lag L): return the values of step s-L, read from the pinned ring the update
        # kernel writes — the host never blocks on the step it has just enqueued (M6: no per-step round trip)
        self._inflight.append(slot)
        if len(self._inflight) > min(lag, self.RING - 2):
            self._last_values = self._wait_slot(self._inflight.popleft())
        # (before the first pipelined step lands these are the values laid down by initialize())
        import dataclasses as _dc
        return _dc.replace(self._last_values, pending=len(self._inflight))

    def _wait_slot(self, slot: int, timeout_s: float = 30.0) -> StepValues:
        """Poll the pinned ring until the device has written step ``_expect[slot]`` (no CUDA call on the way)."""
        import time as _time
        want = self._expect[slot]
        row = self._pinned_u32[slot]  # 16 words: local_step = [2:4]; per-chunk tags (low word of local_step) at 4, 8, 15
        tag = want & 0xFFFFFFFF
        t0 = None
        spins = 0
        while not (int(row[2]) == tag and int(row[3]) == (want >> 32) and int(row[4]) == tag and int(row[8]) == tag
                   and int(row[15]) == tag):  # every 16-byte chunk of this step has landed
            spins += 1
            if spins & 0x3FF == 0:
                now = _time.monotonic()
                t0 = t0 or now
                if now - t0 > timeout_s:
                    break  # fall back to the validated read (synchronises and checks the step counter)
        return self.read_slot(slot)

    def drain(self) -> Optional[StepValues]:
        """Wait for every pipelined step; returns the values of the last one (None when nothing was in flight)."""
        last = None
        while self._inflight:
            last = self._wait_slot(self._inflight.popleft())
        if last is not None:
            self._last_values = last
        return last

    def synchronize(self) -> None:
        self.stream.synchronize()

    def read_values(self) -> StepValues:
        self.drain()
        with torch.cuda.stream(self.stream):
            slot = self.fetch_async()
        self.stream.synchronize()
        return self.read_slot(slot)

    def global_step(self) -> int:
        return self.read_values().global_step

    # ---- checkpoint interface --------------------------------------------------------------------
    def export_variables(self) -> Dict[str, np.ndarray]:
        self.drain()
        with nvtx_range("ttd.checkpoint.export"):
            with torch.cuda.stream(self.stream):
                self.store.snapshot_owner_shards()  # chief-side peer reads; no collective
            self.stream.synchronize()
            return self.store.export_variables()

    def import_variables(self, tensors: Dict[str, np.ndarray]) -> None:
        with torch.cuda.stream(self.stream):
            self.store.import_variables(tensors)
        self.stream.synchronize()

    def close(self) -> None:
        """Peers keep reading / writing this process's memory until THEY are done: leave together."""
        self.stream.synchronize()
        if self.comm is not None and not self._closed:
            self.comm.barrier()
        self._closed = True
        self._graph = None
        self._graph_exec = None
        if self._feeder:
            self.copy_stream.synchronize()
            self.lib.ttd_feeder_destroy(self._feeder)
            self._feeder = None
