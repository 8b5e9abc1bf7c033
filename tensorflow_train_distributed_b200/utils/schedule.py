"""Closed forms of the schedules the kernels evaluate on the device (used by the CPU path and as test oracles)."""
from __future__ import annotations

import math


def exponential_decay(initial_lr: float, global_step: int, decay_steps: int, decay_rate: float,
                      staircase: bool = True) -> float:
    """``tf.train.exponential_decay`` as called at distribute_training.py:136-140:
    ``lr = initial_lr * decay_rate ** (global_step / decay_steps)``, integer division when staircase."""
    p = global_step / float(decay_steps)
    if staircase:
        p = math.floor(p)
    return initial_lr * (decay_rate ** p)


def ema_decay(decay: float, num_updates: int | None) -> float:
    """``tf.train.ExponentialMovingAverage(decay, num_updates)`` (dead code at :153-161 passes global_step):
    effective decay = min(decay, (1 + n) / (10 + n))."""
    if num_updates is None:
        return decay
    return min(decay, (1.0 + num_updates) / (10.0 + num_updates))


def ema_update(shadow: float, value: float, decay: float) -> float:
    """shadow -= (1 - decay) * (shadow - value)"""
    return shadow - (1.0 - decay) * (shadow - value)
