"""``python -m tensorflow_train_distributed_b200.evaluate --checkpoint_dir DIR [--ema] [--data idx --data_path P]``

Loss and top-1 accuracy of a saved model on a data set, dropout off. The reference never evaluates (it hard-codes
``training=True``, distribute_training.py:94, and only prints the training-batch metrics, :232-235); this is the
missing other half for someone who trained with it: restore the TF1 bundle (optionally the moving-average shadows,
:153-161), run the forward pass on the CPU, report. Works for the MLP family and LeNet.
"""
from __future__ import annotations

import sys
from typing import Dict, Optional, Sequence, Tuple

import numpy as np

from .config import TrainConfig, build_parser
from .data import DataSet, make_dataset
from .io import bundle
from .io.inspect import resolve_prefix
from .parallel.sharding import EMA_SUFFIX, ParamLayout


def evaluate(layout: ParamLayout, tensors: Dict[str, np.ndarray], dataset: DataSet, *, activation: str = "elu",
             ema: bool = False, batch_size: int = 1000, max_examples: Optional[int] = None) -> Tuple[float, float, int]:
    """(mean loss, top-1 accuracy, examples) of the variables in ``tensors`` over ``dataset`` in its stored order."""
    import torch

    from .parallel.cpu_ps import CpuMlp
    params = {}
    for spec in layout.variables():
        name = spec.name
        if name not in tensors:
            continue
        src = tensors.get(name + EMA_SUFFIX, tensors[name]) if ema else tensors[name]
        params[name] = torch.from_numpy(np.ascontiguousarray(src, dtype=np.float32))
    model = CpuMlp(layout, activation, 0.0, 0)
    n = dataset.num_examples if max_examples is None else min(max_examples, dataset.num_examples)
    loss_sum, correct = 0.0, 0
    with torch.no_grad():
        for i in range(0, n, batch_size):
            x = torch.from_numpy(dataset.images[i:min(i + batch_size, n)])
            y = torch.from_numpy(dataset.labels[i:min(i + batch_size, n)])
            z = model.logits(params, x, training=False)
            loss_sum += float(torch.nn.functional.cross_entropy(z, y, reduction="sum"))
            correct += int((z.argmax(1) == y).sum())
    return loss_sum / n, correct / n, n


def main(argv: Optional[Sequence[str]] = None) -> int:
    p = build_parser()
    p.prog = "tensorflow_train_distributed_b200.evaluate"
    p.add_argument("--ema", action="store_true", help="evaluate the ExponentialMovingAverage shadows")
    p.add_argument("--max_examples", type=int, default=None)
    ns = p.parse_args(argv)
    ema, max_examples = ns.ema, ns.max_examples
    d = vars(ns)
    d.pop("ema"), d.pop("max_examples")
    cfg = TrainConfig(**d)
    if not cfg.checkpoint_dir:
        p.error("--checkpoint_dir (a directory or a model.ckpt-N prefix) is required")
    from .train import build_layout
    prefix = resolve_prefix(cfg.checkpoint_dir)
    tensors = bundle.read_bundle(prefix)
    dataset = make_dataset(cfg.data, cfg.data_path, cfg.dataset_size, cfg.input_dim, cfg.num_classes, cfg.seed + 1000)
    loss, acc, n = evaluate(build_layout(cfg), tensors, dataset, activation=cfg.activation, ema=ema,
                            max_examples=max_examples)
    print(f"{prefix}: global_step {int(tensors.get('global_step', -1))}, {n} examples, "
          f"loss {loss:.6g}, accuracy {acc:.4f}{' (moving averages)' if ema else ''}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
