"""Micro-benchmark of the fused push (K1) in isolation, under torchrun: device time per call, max over ranks,
for the full kernel and for variants that switch parts of it off (barriers only / no broadcast of the copies)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    from tensorflow_train_distributed_b200.config import TrainConfig
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    from tensorflow_train_distributed_b200.parallel.symmetric import SymmetricComm
    comm = SymmetricComm.from_env()
    rank, world = comm.rank, comm.world_size
    model = os.environ.get("K1_MODEL", "mlp")
    hosts = ",".join(f"127.0.0.1:{2222 + i}" for i in range(world))
    cfg = TrainConfig(worker_hosts=hosts, task_id=rank, sync_replicas=True, save_checkpoint_secs=None, model=model,
                      cuda_graph=False, activation="relu" if model == "wide_mlp" else "elu")
    tr = GpuTrainer(cfg, device_index=comm.device.index, comm=comm)
    tr.initialize()
    st = tr.store
    iters = int(os.environ.get("K1_ITERS", "200"))

    def timeit(label, setup=None):
        """Device time per call with the host out of the picture: 20 calls captured in a CUDA graph."""
        if setup:
            setup()
        comm.barrier()
        per_graph = 20
        with torch.cuda.stream(tr.stream):
            for _ in range(3):
                st.apply_update()
        tr.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=tr.stream):
            for _ in range(per_graph):
                st.apply_update()
        comm.barrier()
        with torch.cuda.stream(tr.stream):
            g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters // per_graph):
                g.replay()
            e1.record()
        tr.synchronize()
        us = comm.max_over_ranks(e0.elapsed_time(e1) * 1000 / (iters // per_graph * per_graph))
        return label, us

    res = {}
    lo, hi, ncopies, grid = st.lo, st.hi, st._peers.n_copies, st._peers.grid
    res.update([timeit("full")])
    trace = st.state.cpu().view(torch.int32)[10:16].tolist()  # TtdStepState.reserved (phase clock of CTA 0, ns)
    res["trace_ns"] = dict(after_barrier_in=trace[0], first_tile_loads=trace[1], tiles_done=trace[2],
                           after_barrier_out=trace[3], tiles=trace[4])

    # slowest CTA of every rank, phase by phase (TtdPeers.reserved bit0), one eager launch
    st._peers.reserved = 1
    comm.barrier()
    with torch.cuda.stream(tr.stream):
        st.state.view(torch.int32)[10:16] = 0
        st.apply_update()
    tr.synchronize()
    mine = st.state.view(torch.int32)[10:15].clone()
    allr = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(allr, mine)
    res["trace_max_over_ctas_ns"] = {f"rank{r}": t.tolist() for r, t in enumerate(allr)}
    st._peers.reserved = 0

    def no_bcast():
        st._peers.n_copies = 1
    res.update([timeit("no_broadcast", no_bcast)])

    def barrier_only():
        st._peers.n_copies = ncopies
        st.hi = st.lo
    res.update([timeit("barriers_only", barrier_only)])

    def small_grid():
        st.hi = hi
        st._peers.grid = max(1, grid // 4)
    res.update([timeit("grid_quarter", small_grid)])

    def one_block_barrier():
        st._peers.grid = 1
        st.hi = st.lo
    res.update([timeit("barriers_only_grid1", one_block_barrier)])
    st._peers.grid = grid
    st.hi = hi
    payload = st.layout.total * 4
    if rank == 0:
        print("K1BENCH " + json.dumps({"world": world, "model": model, "grid": grid, "nvls": bool(st._peers.mc_delta), "params": st.layout.num_parameters,
                                       "us": res, "bytes_grad": payload}), flush=True)
    tr.close()
    dist.barrier(device_ids=[comm.device.index])
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
