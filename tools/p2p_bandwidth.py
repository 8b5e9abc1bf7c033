#!/usr/bin/env python
"""NVLink P2P copy bandwidth between two GPUs of one box (single process, copy engines): the roofline constant
for the push/pull paths. Prints one JSON line."""
import json

import torch


def main() -> None:
    n = torch.cuda.device_count()
    if n < 2:
        print(json.dumps({"p2p": "needs two GPUs"}))
        return
    nbytes = 256 << 20
    a = torch.empty(nbytes, dtype=torch.uint8, device="cuda:0")
    b = torch.empty(nbytes, dtype=torch.uint8, device="cuda:1")
    a2 = torch.empty(nbytes, dtype=torch.uint8, device="cuda:0")
    b2 = torch.empty(nbytes, dtype=torch.uint8, device="cuda:1")
    out = {"bytes": nbytes, "can_access_peer": torch.cuda.can_device_access_peer(0, 1)}
    s0, s1 = torch.cuda.Stream(device=0), torch.cuda.Stream(device=1)

    def run(bidir: bool, reps: int = 10) -> float:
        for _ in range(2):
            with torch.cuda.stream(s0):
                b.copy_(a, non_blocking=True)
            if bidir:
                with torch.cuda.stream(s1):
                    a2.copy_(b2, non_blocking=True)
        torch.cuda.synchronize(0), torch.cuda.synchronize(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(s0):
            e0.record()
            for _ in range(reps):
                b.copy_(a, non_blocking=True)
        if bidir:
            with torch.cuda.stream(s1):
                for _ in range(reps):
                    a2.copy_(b2, non_blocking=True)
        with torch.cuda.stream(s0):
            e1.record()
        torch.cuda.synchronize(0), torch.cuda.synchronize(1)
        return nbytes * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9

    out["unidirectional_GBps"] = run(False)
    out["bidirectional_per_direction_GBps_lower_bound"] = run(True)
    print("P2PBW " + json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
