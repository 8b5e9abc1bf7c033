"""Variable catalogue and sharding plans — the ``replica_device_setter`` of this framework.

The reference wraps graph construction in ``tf.train.replica_device_setter`` (distribute_training.py:186-190):
variables are placed round-robin over the ps tasks in creation order — ``global_step`` first (:116), then
``hiddenK/kernel``, ``hiddenK/bias`` … ``output/kernel``, ``output/bias`` (:53-61) — and every other op stays
on the worker.

Two plans are provided:

``round_robin``  the reference-faithful per-variable placement. Used by the CPU/gloo parameter servers and
                 for checkpoint shard ids.
``flat``         the B200 plan: all trainable variables live in ONE flat fp32 buffer (each dense layer's
                 kernel immediately followed by its bias) cut into N contiguous owner ranges. 85 % of the
                 reference model's parameters are ``hidden1/kernel``, so per-variable placement would leave
                 seven of eight GPUs idle during the update; contiguous ranges balance it and are what a
                 reduce-scatter wants.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Sequence, Tuple

GLOBAL_STEP = "global_step"
EMA_SUFFIX = "/ExponentialMovingAverage"
FLAT_ALIGN = 64  # elements; keeps every group 256-byte aligned


@dataclass(frozen=True)
class VariableSpec:
    name: str
    shape: Tuple[int, ...]
    dtype: str = "float32"
    trainable: bool = True
    regularize: bool = False

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


@dataclass(frozen=True)
class LayerGroup:
    """A dense (or conv-as-matrix) layer: kernel [rows, cols] + bias [cols] stored back to back."""
    name: str                      # e.g. "hidden1"
    rows: int                      # fan-in (for conv: kh*kw*cin)
    cols: int                      # units / output channels
    regularize: bool
    kernel_shape: Tuple[int, ...]  # TF shape of the kernel variable ([in,out] or [kh,kw,cin,cout])
    offset: int = 0                # element offset in the flat buffer

    @property
    def numel(self) -> int:
        return (self.rows + 1) * self.cols

    @property
    def kernel_name(self) -> str:
        return f"{self.name}/kernel"

    @property
    def bias_name(self) -> str:
        return f"{self.name}/bias"


def dense_layer_names(n_layers: int) -> List[str]:
    """hidden1..hidden{n-1}, output (:54, :61)."""
    return [f"hidden{i + 1}" for i in range(n_layers - 1)] + ["output"]


def mlp_groups(input_dim: int, layer_sizes: Sequence[int]) -> List[LayerGroup]:
    """Groups of forward_propagation (:39-63): the regulariser is attached to hidden kernels only (:55 vs :61)."""
    names = dense_layer_names(len(layer_sizes))
    groups, fan_in = [], input_dim
    for i, (name, units) in enumerate(zip(names, layer_sizes)):
        groups.append(LayerGroup(name, fan_in, units, regularize=(i < len(layer_sizes) - 1), kernel_shape=(fan_in, units)))
        fan_in = units
    return groups


def lenet_groups(num_classes: int = 10) -> List[LayerGroup]:
    """LeNet-5 on 28x28x1 with SAME-padded conv1: 61,706 parameters, HWIO conv kernels."""
    return [
        LayerGroup("conv1", 25, 6, True, (5, 5, 1, 6)),
        LayerGroup("conv2", 150, 16, True, (5, 5, 6, 16)),
        LayerGroup("fc1", 400, 120, True, (400, 120)),
        LayerGroup("fc2", 120, 84, True, (120, 84)),
        LayerGroup("output", 84, num_classes, False, (84, num_classes)),
    ]


class ParamLayout:
    """Flat layout of the trainable variables plus the catalogue in TF creation order."""

    def __init__(self, groups: Sequence[LayerGroup]):
        placed, off = [], 0
        for g in groups:
            placed.append(LayerGroup(g.name, g.rows, g.cols, g.regularize, g.kernel_shape, off))
            off += -(-g.numel // FLAT_ALIGN) * FLAT_ALIGN
        self.groups: List[LayerGroup] = placed
        self.total: int = off

    # -- catalogue -----------------------------------------------------------------------------
    def variables(self, with_global_step: bool = True) -> List[VariableSpec]:
        """Creation order of the reference graph: global_step, then kernel/bias per layer."""
        out: List[VariableSpec] = []
        if with_global_step:
            out.append(VariableSpec(GLOBAL_STEP, (), "int64", trainable=False))
        for g in self.groups:
            out.append(VariableSpec(g.kernel_name, tuple(g.kernel_shape), "float32", True, g.regularize))
            out.append(VariableSpec(g.bias_name, (g.cols,), "float32", True, False))
        return out

    @property
    def num_parameters(self) -> int:
        return sum(g.numel for g in self.groups)

    def slices(self) -> Dict[str, Tuple[int, int, Tuple[int, ...]]]:
        """variable name -> (flat offset, numel, TF shape)."""
        out = {}
        for g in self.groups:
            out[g.kernel_name] = (g.offset, g.rows * g.cols, tuple(g.kernel_shape))
            out[g.bias_name] = (g.offset + g.rows * g.cols, g.cols, (g.cols,))
        return out

    def regularize_mask_ranges(self) -> List[Tuple[int, int]]:
        return [(g.offset, g.offset + g.rows * g.cols) for g in self.groups if g.regularize]


# ------------------------------------------------------------------------------------------------
# plans
# ------------------------------------------------------------------------------------------------
def round_robin(variables: Sequence[VariableSpec], n_ps: int) -> Dict[str, int]:
    """``replica_device_setter`` default strategy: variable *i* (creation order) -> ps task ``i mod n_ps``."""
    if n_ps < 1:
        raise ValueError("need at least one ps task")
    return {v.name: i % n_ps for i, v in enumerate(variables)}


@dataclass(frozen=True)
class FlatPlan:
    """Owner ranges over the flat buffer: owner k holds [bounds[k], bounds[k+1])."""
    bounds: Tuple[int, ...]
    n_owners: int

    def range_of(self, owner: int) -> Tuple[int, int]:
        return self.bounds[owner], self.bounds[owner + 1]

    def owner_of(self, flat_index: int) -> int:
        for k in range(self.n_owners):
            if self.bounds[k] <= flat_index < self.bounds[k + 1]:
                return k
        raise IndexError(flat_index)


def flat(layout: ParamLayout, n_owners: int, row_align_first: int = 64) -> FlatPlan:
    """Even contiguous ranges. Boundaries that fall inside the FIRST kernel are snapped to multiples of
    ``row_align_first`` rows so each k-block of the pull-fused first GEMM (K tile = 64 input features) has
    exactly one owner; elsewhere they are snapped to FLAT_ALIGN elements."""
    if n_owners < 1:
        raise ValueError("need at least one owner")
    total = layout.total
    g0 = layout.groups[0]
    first_lo, first_hi = g0.offset, g0.offset + g0.rows * g0.cols
    bounds = [0]
    for k in range(1, n_owners):
        b = (total * k) // n_owners
        if first_lo < b < first_hi and row_align_first > 0:
            unit = row_align_first * g0.cols
            b = first_lo + ((b - first_lo + unit // 2) // unit) * unit
            b = min(b, first_hi)
        else:
            b = (b // FLAT_ALIGN) * FLAT_ALIGN
        bounds.append(max(b, bounds[-1]))
    bounds.append(total)
    return FlatPlan(tuple(bounds), n_owners)


def checkpoint_shard_ids(variables: Sequence[VariableSpec], n_shards: int) -> Dict[str, int]:
    """Shard id recorded for each tensor in a sharded checkpoint = the ps task that owns it."""
    return round_robin(variables, max(n_shards, 1))
