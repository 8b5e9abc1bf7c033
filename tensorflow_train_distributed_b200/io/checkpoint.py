"""Checkpoint directory management: the ``checkpoint`` state file, ``model.ckpt-<step>`` prefixes, keep-last-N.

Counterpart of ``tf.train.Saver(sharded=True, max_to_keep=5)`` + ``CheckpointState`` as used implicitly by
MonitoredTrainingSession (distribute_training.py:209-215). The state file is the same text proto:

    model_checkpoint_path: "model.ckpt-2000"
    all_model_checkpoint_paths: "model.ckpt-0"
    all_model_checkpoint_paths: "model.ckpt-2000"
"""
from __future__ import annotations

import glob
import os
import re
from typing import Dict, List, Mapping, Optional

import numpy as np

from . import bundle

STATE_FILE = "checkpoint"
BASENAME = "model.ckpt"


def _quote(s: str) -> str:
    return '"' + s.replace("\\", "\\\\").replace('"', '\\"') + '"'


def _unquote(s: str) -> str:
    s = s.strip()
    if len(s) >= 2 and s[0] == '"' and s[-1] == '"':
        s = s[1:-1]
    return s.replace('\\"', '"').replace("\\\\", "\\")


def write_state(directory: str, latest: str, all_paths: List[str]) -> None:
    lines = [f"model_checkpoint_path: {_quote(latest)}"] + [f"all_model_checkpoint_paths: {_quote(p)}" for p in all_paths]
    tmp = os.path.join(directory, STATE_FILE + ".tmp")
    with open(tmp, "w") as f:
        f.write("\n".join(lines) + "\n")
    os.replace(tmp, os.path.join(directory, STATE_FILE))


def read_state(directory: str) -> Optional[Dict[str, object]]:
    path = os.path.join(directory, STATE_FILE)
    if not os.path.exists(path):
        return None
    latest, all_paths = None, []
    with open(path) as f:
        for line in f:
            m = re.match(r"\s*(\w+)\s*:\s*(.*)$", line)
            if not m:
                continue
            key, val = m.group(1), _unquote(m.group(2))
            if key == "model_checkpoint_path":
                latest = val
            elif key == "all_model_checkpoint_paths":
                all_paths.append(val)
    if latest is None:
        return None
    return {"model_checkpoint_path": latest, "all_model_checkpoint_paths": all_paths}


def latest_checkpoint(directory: str) -> Optional[str]:
    """Full prefix of the newest checkpoint recorded in ``directory`` (``tf.train.latest_checkpoint``)."""
    st = read_state(directory)
    if st is None:
        return None
    p = st["model_checkpoint_path"]
    prefix = p if os.path.isabs(p) else os.path.join(directory, p)
    return prefix if os.path.exists(prefix + ".index") else None


def checkpoint_step(prefix: str) -> Optional[int]:
    m = re.search(r"-(\d+)$", prefix)
    return int(m.group(1)) if m else None


def _delete_prefix(prefix: str) -> None:
    for f in glob.glob(glob.escape(prefix) + ".index") + glob.glob(glob.escape(prefix) + ".data-*"):
        try:
            os.remove(f)
        except OSError:
            pass


class Saver:
    """Writes ``<dir>/model.ckpt-<global_step>`` bundles, maintains the state file, keeps the newest N."""

    def __init__(self, directory: str, max_to_keep: int = 5, basename: str = BASENAME):
        self.directory, self.max_to_keep, self.basename = directory, max_to_keep, basename
        os.makedirs(directory, exist_ok=True)
        st = read_state(directory)
        self._kept: List[str] = list(st["all_model_checkpoint_paths"]) if st else []

    def prefix_for(self, global_step: int) -> str:
        return os.path.join(self.directory, f"{self.basename}-{int(global_step)}")

    # -- single process --------------------------------------------------------------------------
    def save(self, tensors: Mapping[str, np.ndarray], global_step: int, shard_of: Optional[Mapping[str, int]] = None,
             num_shards: int = 1) -> str:
        prefix = self.prefix_for(global_step)
        bundle.write_bundle(prefix, tensors, shard_of, num_shards)
        self.commit(prefix)
        return prefix

    # -- several shard owners --------------------------------------------------------------------
    def save_shard(self, tensors: Mapping[str, np.ndarray], global_step: int, shard_id: int, num_shards: int) -> str:
        prefix = self.prefix_for(global_step)
        bundle.write_shard(prefix, shard_id, num_shards, tensors)
        return prefix

    def finish(self, global_step: int, num_shards: int) -> str:
        """Chief: merge the shard indices and publish the checkpoint."""
        prefix = self.prefix_for(global_step)
        bundle.merge(prefix, num_shards)
        self.commit(prefix)
        return prefix

    def commit(self, prefix: str) -> None:
        rel = os.path.basename(prefix)
        if rel in self._kept:
            self._kept.remove(rel)
        self._kept.append(rel)
        while self.max_to_keep and len(self._kept) > self.max_to_keep:
            old = self._kept.pop(0)
            _delete_prefix(os.path.join(self.directory, old))
        write_state(self.directory, rel, self._kept)

    def restore_latest(self) -> Optional[Dict[str, np.ndarray]]:
        prefix = latest_checkpoint(self.directory)
        return None if prefix is None else bundle.read_bundle(prefix)
