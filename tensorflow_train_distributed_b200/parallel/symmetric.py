"""Symmetric (peer-mapped) device memory over NVLink — the transport that replaces the reference's gRPC.

The reference moves every parameter and gradient through ``tf.train.Server``'s gRPC channels
(distribute_training.py:175-177, :209). On one 8xB200 box the GPUs can instead map each other's memory:
``torch.distributed._symmetric_memory`` allocates with the CUDA VMM API and exchanges the handles, and hands
back one device pointer per peer (plus an NVSwitch multicast pointer when the fabric supports it). Those raw
pointers are what our own kernels dereference; NCCL is only used here for bootstrap, out-of-band gathers
(checkpoint assembly) and timing reductions.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List

import torch
import torch.distributed as dist


@dataclass
class Arena:
    local: torch.Tensor          # uint8 view of this rank's allocation
    peer_ptrs: List[int]         # device address of every rank's allocation as mapped into THIS process
    multicast_ptr: int           # 0 when NVLS multicast is unavailable
    handle: object


class SymmetricComm:
    def __init__(self, rank: int, world_size: int, device: torch.device):
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed must be initialised (backend nccl) before SymmetricComm")
        self.rank, self.world_size, self.device = rank, world_size, device
        self._arenas: List[Arena] = []

    @classmethod
    def from_env(cls) -> "SymmetricComm":
        """torchrun-style bootstrap: RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT."""
        rank = int(os.environ.get("RANK", "0"))
        world = int(os.environ.get("WORLD_SIZE", "1"))
        local = int(os.environ.get("LOCAL_RANK", str(rank)))
        device = torch.device("cuda", local)
        torch.cuda.set_device(device)
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
        return cls(rank, world, device)

    def alloc(self, nbytes: int) -> Arena:
        import torch.distributed._symmetric_memory as symm_mem
        nbytes = -(-nbytes // 4096) * 4096
        t = symm_mem.empty(nbytes, dtype=torch.uint8, device=self.device)
        hdl = symm_mem.rendezvous(t, group=dist.group.WORLD)
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        mc = 0
        try:
            mc = int(hdl.multicast_ptr or 0)
        except Exception:  # noqa: BLE001 - older handles raise when multicast is unsupported
            mc = 0
        arena = Arena(t, ptrs, mc, hdl)
        self._arenas.append(arena)
        return arena

    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)
        dist.barrier(device_ids=[self.device.index])

    def reset_flags(self, store) -> None:
        """Zero this rank's signal pad (epochs restart with local_step = 0)."""
        store.zero_flags()
        torch.cuda.synchronize(self.device)

    def gather_master(self, store) -> None:
        """Make every rank's fp32 master/EMA complete (owners broadcast their ranges). Checkpoint path only."""
        for owner in range(self.world_size):
            lo, hi = store.plan.range_of(owner)
            if hi > lo:
                dist.broadcast(store.w[lo:hi], src=owner)
                dist.broadcast(store.ema[lo:hi], src=owner)
        torch.cuda.synchronize(self.device)

    def max_over_ranks(self, value: float) -> float:
        t = torch.tensor([value], device=self.device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def close(self) -> None:
        self._arenas.clear()
