"""Flags and hyper-parameters.

The five command-line flags of the reference (``job_name``, ``ps_hosts``, ``worker_hosts``, ``task_id``,
``sync_replicas`` — distribute_training.py:22-36) keep their names, defaults and ``--flag=value`` syntax
(``--task_index`` is accepted as an alias because ``tf.train.Server`` calls it that, :177). Every module
constant (:10-17) and every literal buried in the script (layer sizes :93, dropout / regulariser scale :39,
``save_checkpoint_secs=60`` :213, the 100-step log period :229) is promoted to a flag whose default is the
reference value.
"""
from __future__ import annotations

import argparse
import dataclasses
import datetime as _dt
import os
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

# reference constants, distribute_training.py:10-14
BATCH_SIZE = 128
TRAINING_STEPS = 2000
MOVING_AVERAGE_DECAY = 0.99
LEARNING_RATE_DECAY_FACTOR = 0.96
INITIAL_LEARNING_RATE = 0.01
NUM_EXAMPLES_PER_EPOCH = 60000  # :120 (the schedule assumes 60000 although the TF train split has 55000)

REFERENCE_HIDDEN_UNITS = (200, 100, 50, 25, 10)  # :93
WIDE_HIDDEN_UNITS = (8192, 8192, 10)  # BASELINE.json config 4


def _parse_bool(v) -> bool:
    if isinstance(v, bool):
        return v
    s = str(v).strip().lower()
    if s in ("1", "true", "t", "yes", "y", "on"):
        return True
    if s in ("0", "false", "f", "no", "n", "off"):
        return False
    raise argparse.ArgumentTypeError(f"not a boolean: {v!r}")


def _parse_int_list(v) -> List[int]:
    if isinstance(v, (list, tuple)):
        return [int(x) for x in v]
    return [int(x) for x in str(v).replace(" ", "").split(",") if x != ""]


@dataclass
class TrainConfig:
    # ---- the reference's flags (:24-36) ----
    job_name: str = "worker"
    ps_hosts: str = "localhost:2221"
    worker_hosts: str = "localhost:2222,localhost:2223"
    task_id: int = 0
    sync_replicas: bool = False
    # ---- the reference's constants (:10-17) ----
    batch_size: int = BATCH_SIZE
    training_steps: int = TRAINING_STEPS
    moving_average_decay: float = MOVING_AVERAGE_DECAY
    learning_rate_decay_factor: float = LEARNING_RATE_DECAY_FACTOR
    initial_learning_rate: float = INITIAL_LEARNING_RATE
    num_examples_per_epoch: int = NUM_EXAMPLES_PER_EPOCH
    staircase: bool = True  # :140
    model_save_path: str = "./DistributedModelSave"  # :16 (a Windows path in the reference)
    data_path: str = "./MNIST_Dataset"  # :17
    # ---- literals of the script promoted to flags ----
    model: str = "mlp"  # mlp | wide_mlp | lenet
    hidden_units: List[int] = field(default_factory=lambda: list(REFERENCE_HIDDEN_UNITS))
    input_dim: int = 784  # :191
    num_classes: int = 10
    activation: str = "elu"  # :57 (north-star kernels also provide relu)
    dropout_rate: float = 0.01  # :39
    regularizer: str = "l2"  # none | l1 | l2 — the script builds L1 (:50) that never reaches the loss; north-star asks L2
    regularizer_scale: float = 0.01  # :39
    use_moving_average: bool = True  # north-star; the reference's EMA block is dead code (:153-161)
    replicas_to_aggregate: int = 0  # 0 -> number of workers (:146)
    total_num_replicas: int = 0  # 0 -> number of workers (:147)
    save_checkpoint_secs: Optional[float] = 60.0  # :213; None -> events only (:213-214)
    save_checkpoint_steps: Optional[int] = None
    save_summaries_steps: int = 100  # MonitoredTrainingSession default
    log_step_count_steps: int = 100
    log_every_n_steps: int = 100  # :229
    max_to_keep: int = 5
    checkpoint_dir: Optional[str] = None  # default: <model_save_path>/run-<utc>-checkpoint (:204-205)
    # ---- new-framework switches ----
    device: str = "auto"  # auto | cuda | cpu
    transport: str = "auto"  # auto | nvlink (symmetric memory on GPUs) | gloo (ps processes on CPU, torch.distributed)
    # | tcp (ps processes on CPU: native C++ server on every --ps_hosts endpoint, csrc/io/ps_server.cc)
    sharding: str = "auto"  # auto | flat (GPU store: 1/N ranges of the flat buffer) | round_robin (ps processes)
    seed: int = 1234
    data: str = "synthetic"  # synthetic | idx (MNIST idx files under data_path)
    dataset_size: int = 55000  # size of the TF train split
    prefetch: int = 0  # >0: background reader threads (README "ReadByQueue" variant)
    prefetch_impl: str = "native"  # native: C++ threads filling a pinned ring (csrc/io/loader.cc) | python: queue.Queue
    cuda_graph: bool = True
    pdl: bool = True  # programmatic dependent launch between the kernels of a step
    step_impl: str = "auto"  # auto | column | layers. column: forward + loss + data gradients of a narrow MLP as ONE
    # launch with the activations resident in shared memory (mlp_column_kernel), then every weight gradient — and, with
    # one replica, the optimiser step — as a second launch; the default wherever the model fits. layers: one tcgen05
    # GEMM launch per layer (what wide layers and LeNet's dense head use).
    fused_tail: bool = False  # narrow layers behind the first in ONE launch (measured slower than a launch per
    # layer on B200: one CTA serialises epilogues that separate launches spread over several CTAs; see profiles/)
    tail_impl: str = "chain"  # chain: single-CTA tcgen05 chain kernel | cluster: fp32 CUDA-core cluster kernel |
    # net: EVERY GEMM of the step in one launch of co-resident CTAs with arrival-counter hand-over (gemm_net_kernel)
    pull_fused: bool = False  # sync multi-GPU: fetch W1 from its owners inside the first GEMM (K3) instead of having
    # the update kernel broadcast it. Measured 2-3 us/step SLOWER on the reference MLP (97.8 vs 95.1 us at 8 GPUs: the
    # NVLink fetch lands on the forward's critical path, the broadcast hides behind the barrier), hence off by default.
    fetch_lag: int = 0  # GPU, graph mode: run() returns the values of step s-L, read from the pinned ring the update
    # kernel writes, so the host never waits for the step it just enqueued (0 = the reference's exact per-step fetch;
    # with L > 0 the logged loss / accuracy are L steps old; stop-at-step counts the steps in flight and stays exact)
    write_events: bool = True
    metrics_jsonl: bool = True  # every worker: <checkpoint_dir>/metrics-<task>.jsonl, one line per logging period
    nvtx: bool = False  # NVTX ranges around feed / launch / fetch / checkpoint (Nsight Systems, ncu --nvtx)
    inject_fault: str = ""  # "step=N,rank=R" test hook (kill -> resume)
    rendezvous_timeout_secs: float = 120.0
    barrier_timeout_secs: float = 600.0  # GPU sync replicas: a replica that waits longer than this for its peers inside
    # the update kernel traps (dead-peer detection). Must outlast the chief's checkpoint writes.
    max_restarts: int = 0  # launcher only: relaunch the whole worker group this many times when a worker dies; the
    # new processes resume from the newest checkpoint in --checkpoint_dir (the GPU tier's _RecoverableSession)
    ps_idle_timeout_secs: float = 0.0  # >0: a ps exits after this long without worker traffic (reference: never, :181)

    # ------------------------------------------------------------------ derived values
    @property
    def ps_host_list(self) -> List[str]:
        return [h for h in self.ps_hosts.split(",") if h]

    @property
    def worker_host_list(self) -> List[str]:
        return [h for h in self.worker_hosts.split(",") if h]

    @property
    def n_workers(self) -> int:  # :172
        return len(self.worker_host_list)

    @property
    def n_ps(self) -> int:
        return len(self.ps_host_list)

    @property
    def is_chief(self) -> bool:  # :183
        return self.task_id == 0

    @property
    def decay_steps(self) -> int:  # :120-122
        return int(self.num_examples_per_epoch / self.batch_size)

    @property
    def aggregate_n(self) -> int:
        return self.replicas_to_aggregate or self.n_workers

    @property
    def layer_sizes(self) -> List[int]:
        if self.model == "wide_mlp" and list(self.hidden_units) == list(REFERENCE_HIDDEN_UNITS):
            return list(WIDE_HIDDEN_UNITS)
        return list(self.hidden_units)

    def resolve_checkpoint_dir(self, now: Optional[_dt.datetime] = None) -> str:
        """:204-205 — a per-launch UTC-stamped directory unless --checkpoint_dir pins one (which is what makes
        the auto-resume promised by the comment at :208 actually work)."""
        if self.checkpoint_dir:
            return self.checkpoint_dir
        now = now or _dt.datetime.now(_dt.timezone.utc)
        return os.path.join(self.model_save_path, "run-{}-checkpoint".format(now.strftime("%Y%m%d%H%M%S")))

    def validate(self) -> None:
        if self.job_name not in ("ps", "worker"):
            raise ValueError(f"--job_name must be 'ps' or 'worker', got {self.job_name!r}")
        n = self.n_ps if self.job_name == "ps" else self.n_workers
        if not 0 <= self.task_id < max(n, 1):
            raise ValueError(f"--task_id {self.task_id} out of range for job {self.job_name!r} with {n} tasks")
        if self.activation not in ("elu", "relu", "none"):
            raise ValueError(f"--activation must be elu|relu|none, got {self.activation!r}")
        if self.regularizer not in ("none", "l1", "l2"):
            raise ValueError(f"--regularizer must be none|l1|l2, got {self.regularizer!r}")
        if self.step_impl not in ("auto", "column", "layers"):
            raise ValueError(f"--step_impl must be auto|column|layers, got {self.step_impl!r}")
        if self.model not in ("mlp", "wide_mlp", "lenet"):
            raise ValueError(f"--model must be mlp|wide_mlp|lenet, got {self.model!r}")
        if not 0.0 <= self.dropout_rate < 1.0:
            raise ValueError("--dropout_rate must be in [0, 1)")
        if self.batch_size < 1 or self.training_steps < 0:
            raise ValueError("--batch_size must be positive and --training_steps non-negative")

    def replace(self, **kw) -> "TrainConfig":
        return dataclasses.replace(self, **kw)


class _BoolAction(argparse.Action):
    """TF/absl boolean syntax: ``--flag``, ``--flag=true``, ``--flag false`` and ``--noflag``."""

    def __init__(self, option_strings, dest, **kw):
        opts = list(option_strings)
        for o in option_strings:
            if o.startswith("--"):
                opts.append("--no" + o[2:])
        kw.pop("type", None)
        super().__init__(opts, dest, nargs="?", **kw)

    def __call__(self, parser, namespace, values, option_string=None):
        negated = option_string == "--no" + self.dest
        value = True if values is None else _parse_bool(values)
        setattr(namespace, self.dest, (not value) if negated else value)


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(
        prog="tensorflow_train_distributed_b200",
        description="Parameter-server training (async / sync replicas) on B200 GPUs or CPU processes.",
        allow_abbrev=False,
    )
    d = TrainConfig()
    p.add_argument("--job_name", default=d.job_name, help='"ps" or "worker"')
    p.add_argument("--ps_hosts", default=d.ps_hosts,
                   help='Comma-separated list of hostname:port for the parameter server jobs.')
    p.add_argument("--worker_hosts", default=d.worker_hosts,
                   help='Comma-separated list of hostname:port for the worker jobs.')
    p.add_argument("--task_id", "--task_index", dest="task_id", type=int, default=d.task_id,
                   help="Task ID of the worker/replica running the training.")
    p.add_argument("--sync_replicas", action=_BoolAction, default=d.sync_replicas,
                   help="Aggregate the workers' gradients before applying them (SyncReplicasOptimizer mode).")
    for f in dataclasses.fields(TrainConfig):
        if f.name in ("job_name", "ps_hosts", "worker_hosts", "task_id", "sync_replicas"):
            continue
        default = getattr(d, f.name)
        flag = "--" + f.name
        if f.type in ("bool", bool):
            p.add_argument(flag, action=_BoolAction, default=default)
        elif f.name == "hidden_units":
            p.add_argument(flag, type=_parse_int_list, default=default)
        elif f.type in ("int", int):
            p.add_argument(flag, type=int, default=default)
        elif f.type in ("float", float):
            p.add_argument(flag, type=float, default=default)
        elif f.type in ("Optional[float]",):
            p.add_argument(flag, type=lambda s: None if s.lower() in ("none", "") else float(s), default=default)
        elif f.type in ("Optional[int]",):
            p.add_argument(flag, type=lambda s: None if s.lower() in ("none", "") else int(s), default=default)
        else:
            p.add_argument(flag, type=str, default=default)
    return p


def parse_flags(argv: Optional[Sequence[str]] = None) -> TrainConfig:
    """Equivalent of ``tf.app.run()``'s flag parsing (:241-242)."""
    ns = build_parser().parse_args(argv)
    cfg = TrainConfig(**vars(ns))
    cfg.validate()
    return cfg
