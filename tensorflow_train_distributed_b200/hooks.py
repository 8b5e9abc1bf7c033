"""Session hooks — the ``tf.train.SessionRunHook`` family the reference relies on.

Explicit in the script: ``StopAtStepHook(last_step=TRAINING_STEPS)`` (distribute_training.py:196, :199) and the
sync-replicas hook (:148, :196). Implicit through ``MonitoredTrainingSession`` (:209-215): a
``CheckpointSaverHook`` (save at session start, every 60 s, at end; chief only), a ``StepCounterHook``
(``global_step/sec`` every 100 steps) and a ``SummarySaverHook`` (scalar summaries every 100 steps).
"""
from __future__ import annotations

import os
import signal
import time
from dataclasses import dataclass
from typing import Optional


@dataclass
class SessionRunValues:
    """What ``mon_sess.run([train_op, loss, accuracy, global_step])`` returns (:225)."""
    loss: float
    accuracy: float
    global_step: int
    pending: int = 0  # --fetch_lag: steps already enqueued whose global-step increments are not in global_step yet


@dataclass
class SessionRunContext:
    session: "object"
    local_step: int  # the worker's own step counter (:217, :236)


class SessionRunHook:
    def begin(self) -> None: ...
    def after_create_session(self, session) -> None: ...
    def before_run(self, ctx: SessionRunContext) -> None: ...
    def after_run(self, ctx: SessionRunContext, values: SessionRunValues) -> None: ...
    def end(self, session) -> None: ...


class StopAtStepHook(SessionRunHook):
    """Requests stop once the *global* step reaches ``last_step`` (or ``num_steps`` after the start). In async
    mode the global step is shared, so N workers share the 2000 steps; in sync mode each worker runs ~2000
    (the comment at :220-222)."""

    def __init__(self, num_steps: Optional[int] = None, last_step: Optional[int] = None):
        if (num_steps is None) == (last_step is None):
            raise ValueError("exactly one of num_steps and last_step must be given")
        self._num_steps, self._last_step = num_steps, last_step

    def after_create_session(self, session) -> None:
        if self._last_step is None:
            self._last_step = session.global_step() + self._num_steps
        if session.global_step() >= self._last_step:
            session.request_stop()

    def after_run(self, ctx, values) -> None:
        # every step still in flight adds at least one global step (exactly one with sync replicas), so a pipelined
        # fetch stops on time instead of `fetch_lag` steps late
        if values.global_step + getattr(values, "pending", 0) >= self._last_step:
            ctx.session.request_stop()

    @property
    def last_step(self) -> Optional[int]:
        return self._last_step


class _Timer:
    """SecondOrStepTimer: fires every N seconds or every N steps."""

    def __init__(self, every_secs: Optional[float] = None, every_steps: Optional[int] = None):
        self.every_secs, self.every_steps = every_secs, every_steps
        self._last_time: Optional[float] = None
        self._last_step: Optional[int] = None

    def should_trigger(self, step: int, now: Optional[float] = None) -> bool:
        now = time.time() if now is None else now
        if self._last_step is None:
            return True
        if self._last_step == step:
            return False
        if self.every_secs is not None and now - self._last_time >= self.every_secs:
            return True
        if self.every_steps is not None and step >= self._last_step + self.every_steps:
            return True
        return False

    def update(self, step: int, now: Optional[float] = None):
        now = time.time() if now is None else now
        elapsed = (None, None) if self._last_time is None else (now - self._last_time, step - self._last_step)
        self._last_time, self._last_step = now, step
        return elapsed


class CheckpointSaverHook(SessionRunHook):
    """Saves ``model.ckpt-<global_step>`` at session start, whenever the timer fires, and at the end."""

    def __init__(self, save_secs: Optional[float] = 60.0, save_steps: Optional[int] = None):
        if save_secs is None and save_steps is None:
            raise ValueError("CheckpointSaverHook needs save_secs or save_steps")
        self._timer = _Timer(save_secs, save_steps)
        self.saved_steps = []

    def _save(self, session, step: int) -> None:
        session.save_checkpoint(step)
        self._timer.update(step)
        self.saved_steps.append(step)

    def after_create_session(self, session) -> None:
        self._save(session, session.global_step())

    def after_run(self, ctx, values) -> None:
        if self._timer.should_trigger(values.global_step):
            self._save(ctx.session, values.global_step)

    def end(self, session) -> None:
        step = session.global_step()
        if not self.saved_steps or self.saved_steps[-1] != step:
            self._save(session, step)


class StepCounterHook(SessionRunHook):
    """``global_step/sec`` every N steps (to the events file and kept in ``last_rate``)."""

    def __init__(self, every_n_steps: int = 100, writer=None):
        self._timer = _Timer(every_steps=every_n_steps)
        self._writer = writer
        self.last_rate: Optional[float] = None

    def after_run(self, ctx, values) -> None:
        if self._timer.should_trigger(values.global_step):
            secs, steps = self._timer.update(values.global_step)
            if secs:
                self.last_rate = steps / secs
                if self._writer is not None:
                    self._writer.add_scalar("global_step/sec", self.last_rate, values.global_step)


class SummarySaverHook(SessionRunHook):
    """Scalar summaries named ``loss_<task_id>`` / ``accuracy_<task_id>`` (:128-132), every N steps."""

    def __init__(self, writer, task_id: int = 0, save_steps: int = 100):
        self._writer, self._task, self._timer = writer, task_id, _Timer(every_steps=save_steps)

    def after_run(self, ctx, values) -> None:
        if self._timer.should_trigger(values.global_step):
            self._timer.update(values.global_step)
            self._writer.add_scalar(f"loss_{self._task}", values.loss, values.global_step)
            self._writer.add_scalar(f"accuracy_{self._task}", values.accuracy, values.global_step)
            self._writer.flush()

    def end(self, session) -> None:
        self._writer.flush()


class MetricsJsonlHook(SessionRunHook):
    """One JSON line per logging period — ``{"time", "task", "local_step", "global_step", "loss", "accuracy",
    "steps_per_sec"}`` — in ``<dir>/metrics-<task>.jsonl``. Every worker writes its own file (the events file is
    the chief's, as with MonitoredTrainingSession's SummarySaverHook); machine-readable twin of the console line
    (distribute_training.py:232-235)."""

    def __init__(self, path: str, task_id: int = 0, every_n_steps: int = 100):
        import json as _json
        self._json = _json
        os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
        self._f = open(path, "a", buffering=1)
        self._task, self._every = task_id, max(1, every_n_steps)
        self._t_prev: Optional[float] = None
        self._s_prev = 0

    def after_run(self, ctx, values) -> None:
        step = ctx.local_step + 1
        if step % self._every:
            return
        now = time.time()
        rate = (step - self._s_prev) / (now - self._t_prev) if self._t_prev and now > self._t_prev else None
        self._t_prev, self._s_prev = now, step
        self._f.write(self._json.dumps({"time": now, "task": self._task, "local_step": step,
                                        "global_step": int(values.global_step), "loss": float(values.loss),
                                        "accuracy": float(values.accuracy), "steps_per_sec": rate}) + "\n")

    def end(self, session) -> None:
        self._f.close()


class NanLossHook(SessionRunHook):
    def after_run(self, ctx, values) -> None:
        if values.loss != values.loss:
            raise FloatingPointError(f"loss is NaN at global step {values.global_step}")


class FaultInjectionHook(SessionRunHook):
    """``--inject_fault step=N,rank=R``: kill this process hard when its local step reaches N (exercises the
    kill -> restart -> restore-from-checkpoint path; the reference relies on MonitoredTrainingSession for it, :208)."""

    def __init__(self, spec: str, rank: int):
        kv = dict(p.split("=") for p in spec.split(",") if p)
        self._step, self._rank, self._me = int(kv.get("step", -1)), int(kv.get("rank", 0)), rank

    def after_run(self, ctx, values) -> None:
        if self._me == self._rank and ctx.local_step + 1 == self._step:
            os.kill(os.getpid(), signal.SIGKILL)
