"""``python -m tensorflow_train_distributed_b200.io.inspect <checkpoint dir or prefix> [--tensor NAME] [--to_torch OUT]``

The counterpart of TensorFlow's ``inspect_checkpoint`` for the TF1 tensor bundles this framework writes (and reads):
lists every entry with dtype, shape, data shard and checksum; prints one tensor; or exports the variables as a
PyTorch ``state_dict`` file (kernels transposed from TF's [in, out] to ``nn.Linear``'s [out, in]) — the bridge for
someone moving a reference checkpoint into a stock PyTorch model.
"""
from __future__ import annotations

import argparse
import os
import sys
from typing import Optional, Sequence

import numpy as np

from . import bundle, checkpoint

DTYPE_NAMES = {1: "float32", 2: "float64", 3: "int32", 4: "uint8", 9: "int64", 10: "bool", 14: "bfloat16", 19: "float16"}


def resolve_prefix(path: str) -> str:
    if os.path.isdir(path):
        p = checkpoint.latest_checkpoint(path)
        if p is None:
            raise FileNotFoundError(f"no checkpoint state file in {path}")
        return p
    for suffix in (".index",):
        if path.endswith(suffix):
            return path[:-len(suffix)]
    return path


def describe(prefix: str) -> str:
    lines = []
    with bundle.BundleReader(prefix) as r:
        lines.append(f"{prefix}: {len(r.entries)} tensors in {r.num_shards} data shard(s)")
        total = 0
        for name in r.keys():
            e = r.entries[name]
            n = int(np.prod(e["shape"])) if len(e["shape"]) else 1
            total += n
            lines.append(f"  {name:<48} {DTYPE_NAMES.get(e['dtype'], e['dtype']):<8} {str(tuple(e['shape'])):<18} "
                         f"shard {e['shard_id']} offset {e['offset']} size {e['size']} crc32c(masked) {e['crc32c']:#010x}")
        lines.append(f"  total elements: {total}")
    return "\n".join(lines)


def to_torch_state_dict(prefix: str, ema: bool = False):
    """Variables -> ``{"<layer>.weight": [out, in], "<layer>.bias": [out]}`` (conv kernels HWIO -> OIHW)."""
    import torch
    from ..parallel.sharding import EMA_SUFFIX
    tensors = bundle.read_bundle(prefix)
    out = {}
    for name, v in tensors.items():
        if name.endswith(EMA_SUFFIX) or "/" not in name:
            continue
        src = tensors.get(name + EMA_SUFFIX, v) if ema else v
        layer, kind = name.rsplit("/", 1)
        if kind == "kernel":
            t = torch.from_numpy(np.ascontiguousarray(src))
            out[layer + ".weight"] = (t.permute(3, 2, 0, 1) if t.dim() == 4 else t.t()).contiguous()
        elif kind == "bias":
            out[layer + ".bias"] = torch.from_numpy(np.ascontiguousarray(src))
    if "global_step" in tensors:
        out["global_step"] = torch.tensor(int(tensors["global_step"]))
    return out


def main(argv: Optional[Sequence[str]] = None) -> int:
    ap = argparse.ArgumentParser(prog="tensorflow_train_distributed_b200.io.inspect")
    ap.add_argument("path", help="checkpoint directory (latest checkpoint is used) or a model.ckpt-N prefix")
    ap.add_argument("--tensor", help="print this tensor's values")
    ap.add_argument("--to_torch", help="write a PyTorch state_dict (torch.save) to this file")
    ap.add_argument("--ema", action="store_true", help="--to_torch: export the moving-average shadows instead")
    a = ap.parse_args(argv)
    prefix = resolve_prefix(a.path)
    print(describe(prefix))
    if a.tensor:
        with bundle.BundleReader(prefix) as r:
            np.set_printoptions(threshold=200, linewidth=140)
            print(a.tensor, "=", r.read(a.tensor))
    if a.to_torch:
        import torch
        sd = to_torch_state_dict(prefix, a.ema)
        torch.save(sd, a.to_torch)
        print(f"wrote {len(sd)} entries to {a.to_torch}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
