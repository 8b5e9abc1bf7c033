"""Flags, cluster spec, placement plans, schedules (CPU tier; SURVEY §4 unit row)."""
import datetime as dt
import math

import pytest

from tensorflow_train_distributed_b200 import cluster as cl
from tensorflow_train_distributed_b200.config import TrainConfig, parse_flags
from tensorflow_train_distributed_b200.parallel import sharding
from tensorflow_train_distributed_b200.utils import schedule
from tensorflow_train_distributed_b200.utils import logging as tlog


def test_reference_flag_defaults():
    cfg = parse_flags([])
    assert (cfg.job_name, cfg.ps_hosts, cfg.worker_hosts, cfg.task_id, cfg.sync_replicas) == \
        ("worker", "localhost:2221", "localhost:2222,localhost:2223", 0, False)  # reference :24-36
    assert (cfg.batch_size, cfg.training_steps, cfg.moving_average_decay) == (128, 2000, 0.99)  # :10-12
    assert (cfg.learning_rate_decay_factor, cfg.initial_learning_rate) == (0.96, 0.01)  # :13-14
    assert cfg.hidden_units == [200, 100, 50, 25, 10] and cfg.dropout_rate == 0.01 and cfg.regularizer_scale == 0.01
    assert cfg.decay_steps == 468 and cfg.n_workers == 2 and cfg.n_ps == 1 and cfg.is_chief
    assert cfg.save_checkpoint_secs == 60.0 and cfg.log_every_n_steps == 100


def test_flag_syntax_variants():
    cfg = parse_flags(["--job_name=ps", "--task_index=0", "--sync_replicas", "--ps_hosts", "a:1,b:2"])
    assert cfg.job_name == "ps" and cfg.sync_replicas and cfg.n_ps == 2
    assert parse_flags(["--sync_replicas=True"]).sync_replicas
    assert not parse_flags(["--sync_replicas=false"]).sync_replicas
    assert not parse_flags(["--nosync_replicas"]).sync_replicas
    assert parse_flags(["--task_id", "1"]).task_id == 1
    assert parse_flags(["--hidden_units", "8192,8192,10"]).hidden_units == [8192, 8192, 10]
    assert parse_flags(["--save_checkpoint_secs", "none"]).save_checkpoint_secs is None
    assert parse_flags(["--model", "wide_mlp"]).layer_sizes == [8192, 8192, 10]
    with pytest.raises(SystemExit):
        parse_flags(["--no_such_flag"])
    with pytest.raises(ValueError):
        parse_flags(["--job_name=chief"])
    with pytest.raises(ValueError):
        parse_flags(["--task_id=5"])


def test_checkpoint_dir_naming():
    cfg = TrainConfig(model_save_path="/tmp/m")
    now = dt.datetime(2026, 1, 2, 3, 4, 5, tzinfo=dt.timezone.utc)
    assert cfg.resolve_checkpoint_dir(now) == "/tmp/m/run-20260102030405-checkpoint"  # :204-205, POSIX join
    assert TrainConfig(checkpoint_dir="/x/y").resolve_checkpoint_dir(now) == "/x/y"


def test_cluster_spec_and_rank_maps():
    cfg = TrainConfig(ps_hosts="h:2221,h:2220", worker_hosts="h:2222,h:2223,h:2224")
    c = cl.ClusterSpec.from_config(cfg)
    assert c.as_dict() == {"ps": ["h:2221", "h:2220"], "worker": ["h:2222", "h:2223", "h:2224"]}
    assert c.num_tasks("worker") == 3 and c.task_address("ps", 1) == "h:2220" and c.jobs == ["ps", "worker"]
    t = cl.resolve_task(c, "worker", 0, "gloo")
    assert (t.rank, t.world_size, t.is_chief, t.master_addr, t.master_port) == (2, 5, True, "h", 2221)
    assert cl.resolve_task(c, "ps", 1, "gloo").rank == 1 and not cl.resolve_task(c, "ps", 0, "gloo").is_chief
    g = cl.resolve_task(c, "worker", 2, "nvlink")
    assert (g.rank, g.world_size, g.device_index, g.is_chief, g.master_port) == (2, 3, 2, False, 2222)
    assert cl.resolve_task(c, "ps", 0, "nvlink").rank == -1
    assert cl.worker_ranks(c) == [2, 3, 4] and cl.ps_ranks(c) == [0, 1]
    with pytest.raises(ValueError):
        cl.ClusterSpec({"ps": ["nohostport"]})
    assert cl.choose_transport(TrainConfig(device="cpu"), True) == "tcp"
    assert cl.choose_transport(TrainConfig(), True) == "nvlink" and cl.choose_transport(TrainConfig(), False) == "tcp"


def test_variable_catalogue_and_round_robin():
    layout = sharding.ParamLayout(sharding.mlp_groups(784, [200, 100, 50, 25, 10]))
    names = [v.name for v in layout.variables()]
    assert names == ["global_step", "hidden1/kernel", "hidden1/bias", "hidden2/kernel", "hidden2/bias",
                     "hidden3/kernel", "hidden3/bias", "hidden4/kernel", "hidden4/bias", "output/kernel",
                     "output/bias"]  # creation order, SURVEY A.1
    assert layout.num_parameters == 183685
    v = {x.name: x for x in layout.variables()}
    assert v["hidden1/kernel"].shape == (784, 200) and v["global_step"].dtype == "int64" and not v["global_step"].trainable
    assert v["hidden4/kernel"].regularize and not v["output/kernel"].regularize and not v["hidden1/bias"].regularize
    rr = sharding.round_robin(layout.variables(), 2)
    assert rr["global_step"] == 0 and rr["hidden1/kernel"] == 1 and rr["hidden1/bias"] == 0 and rr["output/bias"] == 0
    assert set(sharding.round_robin(layout.variables(), 1).values()) == {0}
    sl = layout.slices()
    assert sl["hidden1/kernel"][:2] == (0, 156800) and sl["hidden1/bias"][:2] == (156800, 200)
    assert all(g.offset % 64 == 0 for g in layout.groups)


def test_flat_plan_is_balanced_and_k_block_aligned():
    layout = sharding.ParamLayout(sharding.mlp_groups(784, [200, 100, 50, 25, 10]))
    for n in (1, 2, 3, 4, 8):
        p = sharding.flat(layout, n)
        assert p.bounds[0] == 0 and p.bounds[-1] == layout.total and list(p.bounds) == sorted(p.bounds)
        first_end = 784 * 200
        for b in p.bounds[1:-1]:
            if b < first_end:
                assert b % (64 * 200) == 0  # k-blocks of the pull-fused first GEMM have exactly one owner
            else:
                assert b % 64 == 0
        sizes = [p.bounds[i + 1] - p.bounds[i] for i in range(n)]
        assert max(sizes) <= layout.total / n + 64 * 200
        assert p.owner_of(0) == 0 and p.owner_of(layout.total - 1) == n - 1
    wide = sharding.ParamLayout(sharding.mlp_groups(784, [8192, 8192, 10]))
    assert wide.num_parameters == 73629706
    lenet = sharding.ParamLayout(sharding.lenet_groups())
    assert lenet.num_parameters == 61706
    assert [v.shape for v in lenet.variables()[1:3]] == [(5, 5, 1, 6), (6,)]


def test_lr_schedule_and_ema_closed_forms():
    assert schedule.exponential_decay(0.01, 0, 468, 0.96) == 0.01
    assert schedule.exponential_decay(0.01, 467, 468, 0.96) == 0.01
    assert math.isclose(schedule.exponential_decay(0.01, 468, 468, 0.96), 0.0096)
    assert math.isclose(schedule.exponential_decay(0.01, 1999, 468, 0.96), 0.01 * 0.96 ** 4)
    assert math.isclose(schedule.exponential_decay(0.01, 234, 468, 0.96, staircase=False), 0.01 * 0.96 ** 0.5)
    assert schedule.ema_decay(0.99, None) == 0.99
    assert math.isclose(schedule.ema_decay(0.99, 0), 0.1) and math.isclose(schedule.ema_decay(0.99, 80), 0.9)
    assert schedule.ema_decay(0.99, 10 ** 6) == 0.99
    assert math.isclose(schedule.ema_update(1.0, 0.0, 0.9), 0.9)


def test_console_format_matches_reference():
    s = tlog.format_step(100, 201, 0.25, 0.9375, 0.0123)
    assert s == "After 100 training steps (201 global steps), loss on training batch is 0.25, accuracy is 0.9375. (0.012 sec/batch)"
    assert tlog.format_final(1000, 2000) == "total step: 1000, global_step: 2000"


def test_flat_plan_for_every_world_size_and_model():
    """1..8 owners x (reference MLP, wide MLP, LeNet): ranges cover the buffer in order, every owner holds something,
    boundaries are 64-element aligned (the update kernel's 128-bit path needs 4) and balanced to one snap unit."""
    layouts = {"mlp": sharding.ParamLayout(sharding.mlp_groups(784, [200, 100, 50, 25, 10])),
               "wide": sharding.ParamLayout(sharding.mlp_groups(784, [8192, 8192, 10])),
               "lenet": sharding.ParamLayout(sharding.lenet_groups())}
    for name, layout in layouts.items():
        g0 = layout.groups[0]
        unit = max(64 * g0.cols, 64)
        for n in range(1, 9):
            p = sharding.flat(layout, n)
            assert len(p.bounds) == n + 1 and p.bounds[0] == 0 and p.bounds[-1] == layout.total, (name, n)
            sizes = [p.bounds[i + 1] - p.bounds[i] for i in range(n)]
            assert all(s > 0 for s in sizes), (name, n, sizes)
            assert all(b % 4 == 0 for b in p.bounds[1:-1]), (name, n)
            assert max(sizes) - min(sizes) <= 2 * unit + 64, (name, n, sizes)
            assert [p.owner_of(b) for b in p.bounds[:-1]] == list(range(n))
