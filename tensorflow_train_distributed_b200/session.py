"""MonitoredTrainingSession equivalent (distribute_training.py:209-215 and the loop at :223-238).

Responsibilities taken over from TensorFlow:

* chief: restore from the newest checkpoint in ``checkpoint_dir`` if there is one, else initialise; non-chief:
  wait until the chief has done so (``trainer.initialize`` hides the transport-specific handshake);
* run the hook protocol around every step; expose ``should_stop()``;
* chief-only default hooks: checkpoint saver (start / every ``save_checkpoint_secs`` / end), step counter and
  summary saver writing a TensorBoard events file;
* survive a lost parameter server: a ``TransportError`` from the trainer makes the session rebuild itself and
  resume from the newest checkpoint (``_RecoverableSession``).

The trainer is duck-typed — :class:`~.parallel.gpu_trainer.GpuTrainer` (GPU, NVLink) and
:class:`~.parallel.cpu_ps.PsWorkerTrainer` (CPU, gloo) both provide ``initialize / train_step / global_step /
export_variables / close``.
"""
from __future__ import annotations

import os
import time
from typing import Dict, List, Optional, Sequence

import numpy as np

from .hooks import (CheckpointSaverHook, NanLossHook, SessionRunContext, SessionRunHook, SessionRunValues,
                    StepCounterHook, SummarySaverHook)
from .io import checkpoint as ckpt


class TransportError(RuntimeError):
    """The path to the parameter store broke (AbortedError / UnavailableError in TF terms)."""


class MonitoredTrainingSession:
    def __init__(self, trainer, is_chief: bool = True, checkpoint_dir: Optional[str] = None,
                 hooks: Optional[Sequence[SessionRunHook]] = None, save_checkpoint_secs: Optional[float] = 60.0,
                 save_checkpoint_steps: Optional[int] = None, save_summaries_steps: Optional[int] = 100,
                 log_step_count_steps: Optional[int] = 100, max_to_keep: int = 5, task_id: int = 0,
                 num_ckpt_shards: int = 1, shard_of: Optional[Dict[str, int]] = None, write_events: bool = True,
                 max_recoveries: int = 3):
        self.trainer, self.is_chief, self.checkpoint_dir = trainer, is_chief, checkpoint_dir
        self._user_hooks = list(hooks or [])
        self._hooks: List[SessionRunHook] = []
        self._stop = False
        self._local_step = 0
        self._last_values: Optional[SessionRunValues] = None
        self._saver: Optional[ckpt.Saver] = None
        self._events = None
        self._num_shards, self._shard_of = max(1, num_ckpt_shards), shard_of or {}
        self._max_recoveries, self._recoveries = max_recoveries, 0
        self._save_secs, self._save_steps = save_checkpoint_secs, save_checkpoint_steps
        self._summ_steps, self._count_steps = save_summaries_steps, log_step_count_steps
        self._task_id, self._max_to_keep, self._write_events = task_id, max_to_keep, write_events

    # ------------------------------------------------------------------------------------------------
    def __enter__(self) -> "MonitoredTrainingSession":
        restore = None
        if self.checkpoint_dir and self.is_chief:
            self._saver = ckpt.Saver(self.checkpoint_dir, self._max_to_keep)
        if self.checkpoint_dir and os.path.isdir(self.checkpoint_dir):
            prefix = ckpt.latest_checkpoint(self.checkpoint_dir)
            if prefix is not None and (self.is_chief or getattr(self.trainer, "every_rank_restores", False)):
                from .io import bundle
                restore = bundle.read_bundle(prefix)
        self.trainer.initialize(restore=restore)
        self._hooks = list(self._user_hooks) + [NanLossHook()]
        if self.is_chief and self.checkpoint_dir:
            if self._write_events and (self._summ_steps or self._count_steps):
                from .io.events import EventsWriter
                self._events = EventsWriter(self.checkpoint_dir)
            if self._count_steps:
                self._hooks.append(StepCounterHook(self._count_steps, self._events))
            if self._summ_steps and self._events is not None:
                self._hooks.append(SummarySaverHook(self._events, self._task_id, self._summ_steps))
            if self._save_secs is not None or self._save_steps is not None:
                self._hooks.append(CheckpointSaverHook(self._save_secs, self._save_steps))
        for h in self._hooks:
            h.begin()
        for h in self._hooks:
            h.after_create_session(self)
        return self

    def __exit__(self, exc_type, exc, tb) -> bool:
        try:
            if exc_type is None:
                for h in self._hooks:
                    h.end(self)
        finally:
            if self._events is not None:
                self._events.close()
            self.trainer.close()
        return False

    # ------------------------------------------------------------------------------------------------
    def should_stop(self) -> bool:
        return self._stop

    def request_stop(self) -> None:
        self._stop = True

    def global_step(self) -> int:
        return int(self.trainer.global_step())

    @property
    def local_step(self) -> int:
        return self._local_step

    def run(self, x, y) -> SessionRunValues:
        """One ``mon_sess.run([train_op, loss, accuracy, global_step], feed_dict={x: xs, y_: ys})`` (:225-226)."""
        ctx = SessionRunContext(self, self._local_step)
        for h in self._hooks:
            h.before_run(ctx)
        while True:
            try:
                v = self.trainer.train_step(x, y)
                break
            except TransportError:
                self._recover()
        values = SessionRunValues(float(v.loss), float(v.accuracy), int(v.global_step), int(getattr(v, "pending", 0)))
        for h in self._hooks:
            h.after_run(ctx, values)
        self._local_step += 1
        self._last_values = values
        return values

    # ------------------------------------------------------------------------------------------------
    def save_checkpoint(self, global_step: int) -> Optional[str]:
        """Chief: export every variable under the reference's names and write a sharded TF1 bundle."""
        if self._saver is None:
            return None
        tensors = self.trainer.export_variables()
        tensors["global_step"] = np.asarray(int(tensors.get("global_step", global_step)), dtype=np.int64)
        return self._saver.save(tensors, int(tensors["global_step"]), self._shard_of, self._num_shards)

    def _recover(self) -> None:
        """_RecoverableSession: rebuild the connection, restore the newest checkpoint, carry on."""
        self._recoveries += 1
        if self._recoveries > self._max_recoveries:
            raise TransportError(f"giving up after {self._max_recoveries} recoveries")
        time.sleep(min(2.0, 0.2 * self._recoveries))
        restore = None
        if self.checkpoint_dir and (self.is_chief or getattr(self.trainer, "every_rank_restores", False)):
            prefix = ckpt.latest_checkpoint(self.checkpoint_dir)
            if prefix is not None:
                from .io import bundle
                restore = bundle.read_bundle(prefix)
        self.trainer.reconnect(restore=restore)
