"""Command-line entry point — the counterpart of ``python distribute_training.py --job_name=... --task_id=...``.

    python -m tensorflow_train_distributed_b200.train --job_name=ps     --task_id=0 --ps_hosts=... --worker_hosts=...
    python -m tensorflow_train_distributed_b200.train --job_name=worker --task_id=0 --ps_hosts=... --worker_hosts=...

Same control flow as ``main()`` of the reference (distribute_training.py:168-238): build the cluster, ps tasks
serve until the workers are done, workers build model + optimiser, open a monitored session with a
``StopAtStepHook(last_step=TRAINING_STEPS)`` and loop ``while not mon_sess.should_stop()`` feeding batches and
printing the reference's progress line every 100 local steps.
"""
from __future__ import annotations

import os
import sys
import time
from typing import Optional, Sequence

from . import cluster as cl
from .config import TrainConfig, parse_flags
from .data import NativePrefetchQueue, PrefetchQueue, make_dataset
from .hooks import FaultInjectionHook, StopAtStepHook
from .parallel import sharding
from .session import MonitoredTrainingSession
from .utils import logging as tlog


def build_layout(cfg: TrainConfig) -> sharding.ParamLayout:
    if cfg.model == "lenet":
        return sharding.ParamLayout(sharding.lenet_groups(cfg.num_classes))
    return sharding.ParamLayout(sharding.mlp_groups(cfg.input_dim, cfg.layer_sizes))


def run_ps(cfg: TrainConfig, cluster: cl.ClusterSpec, transport: str) -> int:
    """``server.join()`` (:179-181)."""
    task = cl.resolve_task(cluster, "ps", cfg.task_id, transport)
    if transport == "gloo":
        from .parallel.cpu_ps import PsServer, init_gloo
        import torch.distributed as dist
        init_gloo(task, cfg.rendezvous_timeout_secs)
        stats = PsServer(cfg, cluster, task, build_layout(cfg)).serve()
        tlog.log(f"ps {cfg.task_id} done: {stats}")
        dist.destroy_process_group()
        return 0
    if transport == "tcp":
        from .parallel.tcp_ps import NativePsServer
        server = NativePsServer(cfg, cluster, task, build_layout(cfg))
        stats = server.serve()
        server.close()
        tlog.log(f"ps {cfg.task_id} done: {stats}")
        return 0
    # nvlink: the parameter store is GPU memory owned by the workers; a ps process only keeps the CLI contract
    tlog.log(f"ps {cfg.task_id}: parameters live in the workers' GPU memory (nvlink transport); nothing to serve")
    return 0


def run_worker(cfg: TrainConfig, cluster: cl.ClusterSpec, transport: str) -> int:
    task = cl.resolve_task(cluster, "worker", cfg.task_id, transport)
    is_chief = task.is_chief  # :183
    layout = build_layout(cfg)
    # every worker loads the SAME full set (:184) and walks it in its own order
    dataset = make_dataset(cfg.data, cfg.data_path, cfg.dataset_size, cfg.input_dim, cfg.num_classes, cfg.seed + 1000,
                           shuffle_seed=cfg.seed + 1000 * (cfg.task_id + 1) + 1)

    if transport == "gloo":
        from .parallel.cpu_ps import PsWorkerTrainer, init_gloo
        init_gloo(task, cfg.rendezvous_timeout_secs)
        trainer = PsWorkerTrainer(cfg, cluster, task, layout)
        n_shards = cluster.num_tasks("ps")
    elif transport == "tcp":
        from .parallel.cpu_ps import PsWorkerTrainer
        from .parallel.tcp_ps import TcpChannel
        trainer = PsWorkerTrainer(cfg, cluster, task, layout,
                                  channel=TcpChannel(cluster, task.index, cfg.rendezvous_timeout_secs))
        n_shards = cluster.num_tasks("ps")
    else:
        import torch
        from .parallel.gpu_trainer import GpuTrainer
        comm = None
        if task.world_size > 1:
            import torch.distributed as dist
            from .parallel.symmetric import SymmetricComm
            device = torch.device("cuda", task.device_index % torch.cuda.device_count())
            torch.cuda.set_device(device)
            if not dist.is_initialized():
                dist.init_process_group("nccl", init_method=f"tcp://{task.master_addr}:{task.master_port}",
                                        rank=task.rank, world_size=task.world_size, device_id=device)
            comm = SymmetricComm(task.rank, task.world_size, device)
        trainer = GpuTrainer(cfg, device_index=task.device_index if comm is None else comm.device.index, comm=comm)
        n_shards = max(1, cluster.num_tasks("ps"))
    shard_of = sharding.checkpoint_shard_ids(layout.variables(), n_shards)
    for name in list(shard_of):
        shard_of[name + sharding.EMA_SUFFIX] = shard_of[name]

    hooks = [StopAtStepHook(last_step=cfg.training_steps)]  # :196 / :199
    if cfg.inject_fault:
        hooks.append(FaultInjectionHook(cfg.inject_fault, task.index))
    check_point_dir = cfg.resolve_checkpoint_dir()  # :204-205
    if cfg.metrics_jsonl:
        from .hooks import MetricsJsonlHook
        hooks.append(MetricsJsonlHook(os.path.join(check_point_dir, f"metrics-{cfg.task_id}.jsonl"), cfg.task_id,
                                      cfg.log_every_n_steps))
    if cfg.nvtx:
        from .utils import tracing
        tracing.enable_nvtx(True)

    reader = None
    if cfg.prefetch > 0:  # README.md:5-7 "ReadByQueue": reader threads fill a queue; native (C++) unless asked otherwise
        if cfg.prefetch_impl == "native":
            reader = NativePrefetchQueue(dataset, cfg.batch_size, threads=cfg.prefetch, release_lag=cfg.fetch_lag + 3)
        else:
            reader = PrefetchQueue(dataset, cfg.batch_size, threads=cfg.prefetch)
    step, global_step_value = 0, 0
    with MonitoredTrainingSession(trainer, is_chief=is_chief, checkpoint_dir=check_point_dir, hooks=hooks,
                                  save_checkpoint_secs=cfg.save_checkpoint_secs,
                                  save_checkpoint_steps=cfg.save_checkpoint_steps,
                                  save_summaries_steps=cfg.save_summaries_steps,
                                  log_step_count_steps=cfg.log_step_count_steps, max_to_keep=cfg.max_to_keep,
                                  task_id=cfg.task_id, num_ckpt_shards=n_shards, shard_of=shard_of,
                                  write_events=cfg.write_events) as mon_sess:
        tlog.log(tlog.SESSION_STARTED)  # :216
        global_step_value = mon_sess.global_step()
        start_time = time.time()
        while not mon_sess.should_stop():  # :223
            xs, ys = reader.get() if reader is not None else dataset.next_batch(cfg.batch_size)  # :224
            v = mon_sess.run(xs, ys)  # :225-226
            global_step_value = v.global_step
            if step > 0 and step % cfg.log_every_n_steps == 0:  # :229
                duration = time.time() - start_time
                sec_per_batch = duration / max(global_step_value, 1)  # :231 (divides by the GLOBAL step)
                tlog.log(tlog.format_step(step, global_step_value, v.loss, v.accuracy, sec_per_batch))
            step += 1
        drain = getattr(trainer, "drain", None)  # --fetch_lag: steps still in flight
        last = drain() if drain is not None else None
        if last is not None:
            global_step_value = last.global_step
    if reader is not None:
        reader.close()
    tlog.log(tlog.format_final(step, global_step_value))  # :238
    if transport == "gloo":
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    return 0


def main(argv: Optional[Sequence[str]] = None) -> int:
    cfg = parse_flags(argv)
    cluster = cl.ClusterSpec.from_config(cfg)  # :170-173
    cuda = False
    if cfg.device != "cpu":
        try:
            import torch
            cuda = torch.cuda.is_available()
        except Exception:  # noqa: BLE001
            cuda = False
    transport = cl.choose_transport(cfg, cuda)
    # --sharding: the GPU store owns contiguous 1/N ranges of one flat buffer (flat); real ps processes hold whole
    # variables placed round-robin (the reference's replica_device_setter). Asking for the other one is an error.
    want = {"nvlink": "flat"}.get(transport, "round_robin")
    if cfg.sharding not in ("auto", want):
        raise SystemExit(f"--sharding={cfg.sharding} is not available with the {transport} transport (it uses {want})")
    if cfg.job_name == "ps":
        return run_ps(cfg, cluster, transport)
    return run_worker(cfg, cluster, transport)


if __name__ == "__main__":
    sys.exit(main())
