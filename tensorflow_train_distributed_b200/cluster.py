"""Cluster description and role → rank → device mapping.

Counterpart of ``tf.train.ClusterSpec`` + ``tf.train.Server`` bootstrap (distribute_training.py:168-183).
The reference starts one gRPC server per process; here a process only needs to know who it is:

* ``transport == "gloo"``  (CPU, BASELINE config 1): ps tasks are real server processes. Global ranks are
  ``[ps0 .. psP-1, worker0 .. workerW-1]`` and rendezvous happens at the first ps host's address.
* ``transport == "nvlink"`` (GPU): worker *i* drives GPU *i* and *is also* the owner of parameter shard *i*
  (the "ps" is symmetric memory, not a process). ``--job_name=ps`` processes are accepted for CLI
  compatibility and simply wait for the workers to finish, mirroring ``server.join()`` (:179-181).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Tuple

from .config import TrainConfig


class ClusterSpec:
    """``{"ps": [...], "worker": [...]}`` — same constructor shape as tf.train.ClusterSpec (:173)."""

    def __init__(self, jobs: Dict[str, List[str]]):
        self._jobs = {k: list(v) for k, v in jobs.items()}
        for job, hosts in self._jobs.items():
            for h in hosts:
                if ":" not in h:
                    raise ValueError(f"{job} host {h!r} is not hostname:port")

    @classmethod
    def from_config(cls, cfg: TrainConfig) -> "ClusterSpec":
        return cls({"ps": cfg.ps_host_list, "worker": cfg.worker_host_list})  # :170-173

    @property
    def jobs(self) -> List[str]:
        return sorted(self._jobs)

    def num_tasks(self, job: str) -> int:
        return len(self._jobs.get(job, []))

    def job_tasks(self, job: str) -> List[str]:
        return list(self._jobs.get(job, []))

    def task_address(self, job: str, index: int) -> str:
        return self._jobs[job][index]

    def as_dict(self) -> Dict[str, List[str]]:
        return {k: list(v) for k, v in self._jobs.items()}

    def __eq__(self, other) -> bool:
        return isinstance(other, ClusterSpec) and self._jobs == other._jobs

    def __repr__(self) -> str:
        return f"ClusterSpec({self._jobs!r})"


@dataclass(frozen=True)
class TaskInfo:
    job: str
    index: int
    rank: int          # global rank in the process group of the chosen transport (-1: not a member)
    world_size: int
    is_chief: bool     # :183 — worker task 0
    device_index: int  # GPU ordinal for nvlink workers, -1 otherwise
    master_addr: str
    master_port: int


def _split(addr: str) -> Tuple[str, int]:
    host, port = addr.rsplit(":", 1)
    return host, int(port)


def resolve_task(cluster: ClusterSpec, job: str, index: int, transport: str) -> TaskInfo:
    """Map (job, task index) to a rank of the transport's process group."""
    n_ps, n_w = cluster.num_tasks("ps"), cluster.num_tasks("worker")
    if job not in ("ps", "worker"):
        raise ValueError(f"unknown job {job!r}")
    if not 0 <= index < cluster.num_tasks(job):
        raise ValueError(f"task index {index} out of range for job {job!r}")
    chief = job == "worker" and index == 0
    if transport == "gloo":
        if n_ps < 1:
            raise ValueError("the gloo transport needs at least one ps task")
        host, port = _split(cluster.task_address("ps", 0))
        rank = index if job == "ps" else n_ps + index
        return TaskInfo(job, index, rank, n_ps + n_w, chief, -1, host, port)
    if transport == "tcp":  # no process group: every ps listens on its own --ps_hosts entry (:173-177)
        if n_ps < 1:
            raise ValueError("the tcp transport needs at least one ps task")
        host, port = _split(cluster.task_address("ps", 0))
        return TaskInfo(job, index, index if job == "ps" else n_ps + index, n_ps + n_w, chief, -1, host, port)
    if transport == "nvlink":
        host, port = _split(cluster.task_address("worker", 0))
        if job == "ps":
            return TaskInfo(job, index, -1, n_w, False, -1, host, port)
        return TaskInfo(job, index, index, n_w, chief, index, host, port)
    raise ValueError(f"unknown transport {transport!r}")


def ps_ranks(cluster: ClusterSpec) -> List[int]:
    return list(range(cluster.num_tasks("ps")))


def worker_ranks(cluster: ClusterSpec) -> List[int]:
    n_ps = cluster.num_tasks("ps")
    return [n_ps + i for i in range(cluster.num_tasks("worker"))]


def choose_transport(cfg: TrainConfig, cuda_available: bool) -> str:
    if cfg.transport != "auto":
        return cfg.transport
    # CPU tier: the native TCP parameter server (measured 2-3x the gloo tier's steps/s);
    # --transport gloo keeps the torch.distributed variant
    if cfg.device == "cpu":
        return "tcp"
    if cfg.device == "cuda" or cuda_available:
        return "nvlink"
    return "tcp"
