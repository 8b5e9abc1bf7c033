"""NVLink flag round-trip latency with different ordering recipes (run under torchrun with 2 ranks)."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    from tensorflow_train_distributed_b200.ops import _native as N
    from tensorflow_train_distributed_b200.parallel.symmetric import SymmetricComm
    comm = SymmetricComm.from_env()
    assert comm.world_size == 2
    lib = N.load()
    lib.ttd_diag_pingpong.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    arena = comm.alloc(4096)
    out = torch.zeros(1, dtype=torch.int64, device=comm.device)
    res = {}
    iters = 2000
    for variant, name in enumerate(["volatile", "fence+relaxed", "release/acquire", "atomicExch", "red.release"]):
        arena.local.zero_()
        comm.barrier()
        mine = arena.peer_ptrs[comm.rank] + 256 * variant
        theirs = arena.peer_ptrs[1 - comm.rank] + 256 * variant
        N.check(lib.ttd_diag_pingpong(mine, theirs, comm.rank, iters, variant, out.data_ptr(), N.cur_stream()), "pp")
        torch.cuda.synchronize()
        res[name] = out.item() / iters / 1000.0
        comm.barrier()
    if comm.rank == 0:
        print("PINGPONG_RTT_US " + json.dumps(res), flush=True)
    dist.barrier(device_ids=[comm.device.index])
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
