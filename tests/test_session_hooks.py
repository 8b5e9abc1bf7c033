"""Hook protocol and the monitored session against a fake trainer (CPU tier)."""
import os

import numpy as np
import pytest

from tensorflow_train_distributed_b200 import hooks as H
from tensorflow_train_distributed_b200.io import bundle, checkpoint
from tensorflow_train_distributed_b200.session import MonitoredTrainingSession, TransportError


class FakeTrainer:
    def __init__(self, fail_at=None, step_inc=1):
        self.gs, self.init_restore, self.closed, self.fail_at, self.inc, self.reconnects = 0, None, False, fail_at, step_inc, 0

    def initialize(self, restore=None):
        self.init_restore = restore
        if restore is not None:
            self.gs = int(restore["global_step"])

    def reconnect(self, restore=None):
        self.reconnects += 1
        self.initialize(restore)

    def train_step(self, x, y):
        if self.fail_at is not None and self.gs == self.fail_at:
            self.fail_at = None
            raise TransportError("ps went away")
        self.gs += self.inc
        return H.SessionRunValues(1.0 / self.gs, 0.5, self.gs)

    def global_step(self):
        return self.gs

    def export_variables(self):
        return {"global_step": np.asarray(self.gs, dtype=np.int64), "w": np.full((3,), float(self.gs), dtype=np.float32)}

    def close(self):
        self.closed = True


def test_stop_at_step_on_global_step():
    tr = FakeTrainer(step_inc=2)  # another worker advances the shared global step too (async mode)
    with MonitoredTrainingSession(tr, hooks=[H.StopAtStepHook(last_step=10)]) as s:
        n = 0
        while not s.should_stop():
            s.run(None, None)
            n += 1
    assert n == 5 and tr.gs == 10 and tr.closed and s.local_step == 5
    with pytest.raises(ValueError):
        H.StopAtStepHook()
    tr2 = FakeTrainer()
    tr2.gs = 50
    with MonitoredTrainingSession(tr2, hooks=[H.StopAtStepHook(num_steps=3)]) as s:
        while not s.should_stop():
            s.run(None, None)
    assert tr2.gs == 53


def test_checkpoint_hook_saves_start_timer_end_and_resume(tmp_path):
    d = str(tmp_path / "run")
    tr = FakeTrainer()
    with MonitoredTrainingSession(tr, is_chief=True, checkpoint_dir=d, hooks=[H.StopAtStepHook(last_step=7)],
                                  save_checkpoint_secs=None, save_checkpoint_steps=3) as s:
        while not s.should_stop():
            s.run(None, None)
    st = checkpoint.read_state(d)
    assert st["all_model_checkpoint_paths"] == ["model.ckpt-0", "model.ckpt-3", "model.ckpt-6", "model.ckpt-7"]
    assert any(f.startswith("events.out.tfevents.") for f in os.listdir(d))
    # a restarted chief pointed at the same directory resumes (what the comment at :208 promises)
    tr2 = FakeTrainer()
    with MonitoredTrainingSession(tr2, is_chief=True, checkpoint_dir=d, hooks=[H.StopAtStepHook(last_step=9)],
                                  save_checkpoint_secs=None, save_checkpoint_steps=100) as s:
        assert tr2.init_restore is not None and s.global_step() == 7
        while not s.should_stop():
            s.run(None, None)
    assert tr2.gs == 9
    assert int(bundle.read_bundle(checkpoint.latest_checkpoint(d))["global_step"]) == 9
    # non-chief never writes
    d2 = str(tmp_path / "other")
    with MonitoredTrainingSession(FakeTrainer(), is_chief=False, checkpoint_dir=d2, hooks=[H.StopAtStepHook(last_step=2)]) as s:
        while not s.should_stop():
            s.run(None, None)
    assert not os.path.exists(d2)


def test_timer_semantics():
    t = H._Timer(every_secs=60)
    assert t.should_trigger(0, now=0.0)
    t.update(0, now=0.0)
    assert not t.should_trigger(5, now=59.0) and t.should_trigger(5, now=60.0) and not t.should_trigger(0, now=100.0)
    t2 = H._Timer(every_steps=100)
    t2.update(0, now=0.0)
    assert not t2.should_trigger(99) and t2.should_trigger(100)
    assert t2.update(100, now=4.0) == (4.0, 100)


def test_summaries_and_step_counter(tmp_path):
    from tensorflow_train_distributed_b200.io import events
    d = str(tmp_path / "ev")
    with MonitoredTrainingSession(FakeTrainer(), is_chief=True, checkpoint_dir=d, task_id=3,
                                  hooks=[H.StopAtStepHook(last_step=250)], save_checkpoint_secs=None,
                                  save_summaries_steps=100, log_step_count_steps=100) as s:
        while not s.should_stop():
            s.run(None, None)
    f = [os.path.join(d, x) for x in os.listdir(d) if x.startswith("events")][0]
    tags = [t for e in events.read_events(f) for t in e.get("scalars", {})]
    assert tags.count("loss_3") == 3 and tags.count("accuracy_3") == 3 and "global_step/sec" in tags  # :128-132


def test_recoverable_session_restores_latest_checkpoint(tmp_path):
    d = str(tmp_path / "rec")
    tr = FakeTrainer(fail_at=5)
    with MonitoredTrainingSession(tr, is_chief=True, checkpoint_dir=d, hooks=[H.StopAtStepHook(last_step=8)],
                                  save_checkpoint_secs=None, save_checkpoint_steps=2) as s:
        while not s.should_stop():
            s.run(None, None)
    assert tr.reconnects == 1 and tr.gs == 8
    assert int(tr.init_restore["global_step"]) == 4  # resumed from model.ckpt-4, redid step 5


def test_nan_loss_raises():
    class Nan(FakeTrainer):
        def train_step(self, x, y):
            self.gs += 1
            return H.SessionRunValues(float("nan"), 0.0, self.gs)
    with pytest.raises(FloatingPointError):
        with MonitoredTrainingSession(Nan()) as s:
            s.run(None, None)


def test_metrics_jsonl_hook_writes_one_line_per_period(tmp_path):
    import json
    from tensorflow_train_distributed_b200.hooks import MetricsJsonlHook, SessionRunContext, SessionRunValues
    path = str(tmp_path / "run" / "metrics-1.jsonl")
    h = MetricsJsonlHook(path, task_id=1, every_n_steps=10)
    for step in range(35):
        h.after_run(SessionRunContext(session=None, local_step=step), SessionRunValues(1.0 / (step + 1), 0.5, 2 * step))
    h.end(None)
    rows = [json.loads(line) for line in open(path)]
    assert [r["local_step"] for r in rows] == [10, 20, 30]
    assert rows[0]["task"] == 1 and rows[1]["global_step"] == 38 and rows[0]["steps_per_sec"] is None
    assert rows[2]["steps_per_sec"] > 0


def test_nvtx_range_is_a_noop_without_cuda():
    from tensorflow_train_distributed_b200.utils import tracing
    tracing.enable_nvtx(True)
    with tracing.nvtx_range("x"):
        pass
    tracing.enable_nvtx(False)
    assert not tracing.nvtx_enabled()


def test_stop_at_step_counts_steps_in_flight():
    """--fetch_lag: values lag, but every enqueued step will add >= 1 global step, so the stop comes on time."""
    from tensorflow_train_distributed_b200.hooks import SessionRunContext, SessionRunValues, StopAtStepHook

    class _S:
        stopped = False

        def global_step(self):
            return 0

        def request_stop(self):
            self.stopped = True

    s = _S()
    h = StopAtStepHook(last_step=10)
    h.after_create_session(s)
    h.after_run(SessionRunContext(s, 0), SessionRunValues(1.0, 0.5, 5, pending=4))
    assert not s.stopped
    h.after_run(SessionRunContext(s, 1), SessionRunValues(1.0, 0.5, 6, pending=4))
    assert s.stopped
