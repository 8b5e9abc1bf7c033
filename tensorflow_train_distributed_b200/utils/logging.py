"""Console output in the reference's exact wording (distribute_training.py:216, :229-238)."""
from __future__ import annotations

import sys

SESSION_STARTED = "session started"  # :216
STEP_FORMAT = ("After %d training steps (%d global steps), "
               "loss on training batch is %g, "
               "accuracy is %g. (%.3f sec/batch)")  # :232-234
FINAL_FORMAT = "total step: %d, global_step: %d"  # :238


def format_step(step: int, global_step: int, loss: float, accuracy: float, sec_per_batch: float) -> str:
    return STEP_FORMAT % (step, global_step, loss, accuracy, sec_per_batch)


def format_final(step: int, global_step: int) -> str:
    return FINAL_FORMAT % (step, global_step)


def log(msg: str, file=None) -> None:
    print(msg, file=file or sys.stdout, flush=True)
