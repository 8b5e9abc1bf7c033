"""TF1 tensor-bundle writer/reader, crc32c, SSTable format and checkpoint state (CPU tier)."""
import os
import struct

import numpy as np
import pytest

from tensorflow_train_distributed_b200.io import bundle, checkpoint, events


def test_crc32c_known_vectors_and_paths_agree():
    # RFC 3720 test vectors
    assert bundle.crc32c(b"") == 0
    assert bundle.crc32c(b"123456789") == 0xE3069283
    assert bundle.crc32c(bytes(32)) == 0x8A9136AA
    assert bundle.crc32c(bytes([0xFF] * 32)) == 0x62A8AB43
    assert bundle.crc32c(bytes(range(32))) == 0x46DD794E
    rng = np.random.default_rng(0)
    for n in (1, 7, 8, 9, 63, 64, 65, 1000, 4097):
        b = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert bundle.crc32c(b) == bundle.crc32c_software(b)


def test_crc_mask_formula():
    c = bundle.crc32c(b"123456789")
    want = (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF
    assert bundle.masked_crc32c(b"123456789") == want


def _ref_tensors(seed=0):
    rng = np.random.default_rng(seed)
    t = {"global_step": np.asarray(1234, dtype=np.int64)}
    for name, shape in [("hidden1/kernel", (784, 200)), ("hidden1/bias", (200,)), ("hidden2/kernel", (200, 100)),
                        ("hidden2/bias", (100,)), ("hidden3/kernel", (100, 50)), ("hidden3/bias", (50,)),
                        ("hidden4/kernel", (50, 25)), ("hidden4/bias", (25,)), ("output/kernel", (25, 10)),
                        ("output/bias", (10,))]:
        t[name] = rng.standard_normal(shape).astype(np.float32)
    return t


def test_single_shard_roundtrip(tmp_path):
    t = _ref_tensors()
    prefix = str(tmp_path / "model.ckpt-1234")
    bundle.write_bundle(prefix, t)
    assert sorted(os.listdir(tmp_path)) == ["model.ckpt-1234.data-00000-of-00001", "model.ckpt-1234.index"]
    with bundle.BundleReader(prefix) as r:
        assert r.num_shards == 1
        assert r.keys() == sorted(t)  # 11 entries for the async reference graph
        assert len(r.keys()) == 11
        assert r.entries["global_step"]["dtype"] == bundle.DT_INT64 and r.entries["global_step"]["shape"] == ()
        assert r.entries["hidden1/kernel"]["shape"] == (784, 200)
        for k, v in t.items():
            got = r.read(k)
            assert got.dtype == v.dtype and got.shape == v.shape and np.array_equal(got, v)
    # data shard = tensors concatenated in key order, no padding
    blob = open(prefix + ".data-00000-of-00001", "rb").read()
    assert blob == b"".join(np.ascontiguousarray(t[k]).tobytes() for k in sorted(t))


def test_sharded_roundtrip_round_robin(tmp_path):
    from tensorflow_train_distributed_b200.parallel import sharding
    layout = sharding.ParamLayout(sharding.mlp_groups(784, [200, 100, 50, 25, 10]))
    shard_of = sharding.round_robin(layout.variables(), 3)
    assert shard_of["global_step"] == 0 and shard_of["hidden1/kernel"] == 1 and shard_of["hidden1/bias"] == 2
    t = _ref_tensors(1)
    prefix = str(tmp_path / "model.ckpt-7")
    for k in range(3):  # what three ps owners would each do
        bundle.write_shard(prefix, k, 3, {n: v for n, v in t.items() if shard_of[n] == k})
    assert os.path.isdir(prefix + "_temp")
    bundle.merge(prefix, 3)
    assert not os.path.exists(prefix + "_temp")
    names = sorted(os.listdir(tmp_path))
    assert names == ["model.ckpt-7.data-00000-of-00003", "model.ckpt-7.data-00001-of-00003",
                     "model.ckpt-7.data-00002-of-00003", "model.ckpt-7.index"]
    with bundle.BundleReader(prefix) as r:
        assert r.num_shards == 3
        for k, v in t.items():
            assert r.entries[k]["shard_id"] == shard_of[k]
            assert np.array_equal(r.read(k), v)


def test_index_file_layout_golden(tmp_path):
    """Byte-level checks of the SSTable framing against the format spec."""
    prefix = str(tmp_path / "g")
    bundle.write_bundle(prefix, {"a": np.asarray([1.0, 2.0], dtype=np.float32), "b": np.asarray(5, dtype=np.int64)})
    raw = open(prefix + ".index", "rb").read()
    assert len(raw) >= 48
    assert struct.unpack("<Q", raw[-8:])[0] == 0xDB4775248B80FB57          # table magic
    footer = raw[-48:]
    # metaindex handle then index handle (varint offset, varint size), zero padding up to 40 bytes
    def varint(buf, pos):
        out = shift = 0
        while True:
            b = buf[pos]; pos += 1
            out |= (b & 0x7F) << shift
            if not b & 0x80:
                return out, pos
            shift += 7
    moff, p = varint(footer, 0); msize, p = varint(footer, p); ioff, p = varint(footer, p); isize, p = varint(footer, p)
    assert footer[p:40] == bytes(40 - p)
    assert raw[moff:moff + msize] == bytes([0, 0, 0, 0, 1, 0, 0, 0])       # empty metaindex block
    assert moff + msize + 5 == ioff                                          # 5-byte trailer between blocks
    # first data block starts at 0; its first entry is the header under the empty key
    assert raw[0:3] == bytes([0, 0, 6])                                      # shared=0 non_shared=0 value_len=6
    assert raw[3:9] == bytes([0x08, 0x01, 0x1A, 0x02, 0x08, 0x01])          # num_shards=1, version{producer=1}
    # entry "a": dtype float, shape [2], size 8, crc
    a_bytes = np.asarray([1.0, 2.0], dtype=np.float32).tobytes()
    want_entry = bytes([0x08, 0x01, 0x12, 0x04, 0x12, 0x02, 0x08, 0x02, 0x28, 0x08, 0x35]) + struct.pack("<I", bundle.masked_crc32c(a_bytes))
    assert raw[9:12] == bytes([0, 1, len(want_entry)]) and raw[12:13] == b"a"
    assert raw[13:13 + len(want_entry)] == want_entry
    # block trailer: type 0 + masked crc over contents||type
    data_block_end = moff - 5
    assert raw[data_block_end] == 0
    assert struct.unpack("<I", raw[data_block_end + 1:data_block_end + 5])[0] == bundle.masked_crc32c(raw[:data_block_end + 1])


def test_many_entries_span_restart_points_and_blocks(tmp_path):
    t = {f"layer{i:04d}/w": np.full((3,), i, dtype=np.float32) for i in range(5000)}
    prefix = str(tmp_path / "big")
    bundle.write_bundle(prefix, t)
    got = bundle.read_bundle(prefix)
    assert len(got) == 5000 and all(np.array_equal(got[k], t[k]) for k in ("layer0000/w", "layer2500/w", "layer4999/w"))


def test_dtypes_and_empty_tensor(tmp_path):
    t = {"f64": np.arange(6, dtype=np.float64).reshape(2, 3), "i32": np.arange(4, dtype=np.int32),
         "u8": np.arange(5, dtype=np.uint8), "b": np.asarray([True, False]), "h": np.arange(3, dtype=np.float16),
         "empty": np.zeros((0, 4), dtype=np.float32)}
    prefix = str(tmp_path / "d")
    bundle.write_bundle(prefix, t)
    got = bundle.read_bundle(prefix)
    for k, v in t.items():
        assert got[k].dtype == v.dtype and got[k].shape == v.shape and np.array_equal(got[k], v)


def test_corruption_is_detected(tmp_path):
    prefix = str(tmp_path / "c")
    bundle.write_bundle(prefix, _ref_tensors())
    data = prefix + ".data-00000-of-00001"
    raw = bytearray(open(data, "rb").read())
    raw[100] ^= 0xFF
    open(data, "wb").write(bytes(raw))
    with bundle.BundleReader(prefix) as r:
        with pytest.raises(bundle.BundleError, match="checksum"):
            r.read_all()
    idx = bytearray(open(prefix + ".index", "rb").read())
    idx[5] ^= 0x01
    open(prefix + ".index", "wb").write(bytes(idx))
    with pytest.raises(bundle.BundleError, match="checksum"):
        bundle.BundleReader(prefix)
    with pytest.raises(bundle.BundleError):
        bundle.BundleReader(str(tmp_path / "missing"))


def test_checkpoint_state_and_keep_last_n(tmp_path):
    d = str(tmp_path / "ck")
    s = checkpoint.Saver(d, max_to_keep=3)
    assert checkpoint.latest_checkpoint(d) is None
    for step in (0, 100, 200, 300, 400):
        s.save({"global_step": np.asarray(step, dtype=np.int64), "w": np.full((4,), step, dtype=np.float32)}, step)
    st = checkpoint.read_state(d)
    assert st["model_checkpoint_path"] == "model.ckpt-400"
    assert st["all_model_checkpoint_paths"] == ["model.ckpt-200", "model.ckpt-300", "model.ckpt-400"]
    text = open(os.path.join(d, "checkpoint")).read()
    assert text.splitlines()[0] == 'model_checkpoint_path: "model.ckpt-400"'
    assert not os.path.exists(os.path.join(d, "model.ckpt-0.index")) and os.path.exists(os.path.join(d, "model.ckpt-200.index"))
    assert checkpoint.latest_checkpoint(d) == os.path.join(d, "model.ckpt-400")
    assert checkpoint.checkpoint_step(checkpoint.latest_checkpoint(d)) == 400
    got = checkpoint.Saver(d).restore_latest()
    assert int(got["global_step"]) == 400 and np.array_equal(got["w"], np.full((4,), 400, dtype=np.float32))


def test_events_writer_roundtrip(tmp_path):
    w = events.EventsWriter(str(tmp_path))
    assert os.path.basename(w.path).startswith("events.out.tfevents.")
    w.add_scalar("loss_0", 0.5, 100, wall_time=12.5)
    w.add_scalar("accuracy_0", 0.875, 100)
    w.add_scalar("global_step/sec", 2500.0, 200)
    w.close()
    evs = events.read_events(w.path)
    assert evs[0]["file_version"] == "brain.Event:2"
    assert evs[1]["step"] == 100 and evs[1]["wall_time"] == 12.5 and evs[1]["scalars"] == {"loss_0": 0.5}
    assert evs[2]["scalars"] == {"accuracy_0": 0.875} and evs[3]["scalars"] == {"global_step/sec": 2500.0}


def test_inspect_cli_lists_and_exports_a_torch_state_dict(tmp_path, capsys):
    import torch
    from tensorflow_train_distributed_b200.io import checkpoint, inspect as insp
    rng = np.random.default_rng(0)
    tensors = {"global_step": np.asarray(7, dtype=np.int64),
               "hidden1/kernel": rng.standard_normal((784, 200)).astype(np.float32), "hidden1/bias": np.zeros(200, np.float32),
               "conv1/kernel": rng.standard_normal((5, 5, 1, 6)).astype(np.float32), "conv1/bias": np.ones(6, np.float32),
               "hidden1/kernel/ExponentialMovingAverage": np.full((784, 200), 0.5, np.float32)}
    saver = checkpoint.Saver(str(tmp_path))
    saver.save(tensors, 7)
    out_file = str(tmp_path / "sd.pt")
    assert insp.main([str(tmp_path), "--tensor", "conv1/bias", "--to_torch", out_file]) == 0
    text = capsys.readouterr().out
    assert "6 tensors in 1 data shard" in text and "hidden1/kernel" in text and "(784, 200)" in text
    assert "conv1/bias = [1. 1. 1. 1. 1. 1.]" in text
    sd = torch.load(out_file)
    assert sd["hidden1.weight"].shape == (200, 784) and int(sd["global_step"]) == 7
    x = torch.randn(3, 784)
    assert torch.allclose(torch.nn.functional.linear(x, sd["hidden1.weight"], sd["hidden1.bias"]),
                          x @ torch.from_numpy(tensors["hidden1/kernel"]), atol=1e-3, rtol=1e-4)
    assert sd["conv1.weight"].shape == (6, 1, 5, 5)
    assert torch.equal(sd["conv1.weight"][2, 0], torch.from_numpy(tensors["conv1/kernel"][:, :, 0, 2]))
    ema = insp.to_torch_state_dict(insp.resolve_prefix(str(tmp_path)), ema=True)
    assert float(ema["hidden1.weight"].mean()) == 0.5


def test_reader_survives_corrupted_index_files(tmp_path):
    """Truncations, bit flips, 0xff runs and insertions in the .index: the native reader reports an error (or
    returns checksummed tensors) — it never crashes and never returns silently wrong bytes."""
    import random
    import shutil
    from tensorflow_train_distributed_b200.io import bundle
    rng = np.random.default_rng(0)
    tensors = {f"layer{i}/kernel": rng.standard_normal((7 + i, 5)).astype(np.float32) for i in range(40)}
    prefix = str(tmp_path / "model.ckpt-3")
    bundle.write_bundle(prefix, tensors)
    idx = open(prefix + ".index", "rb").read()
    rnd = random.Random(1234)
    opened = 0
    for it in range(150):
        b = bytearray(idx)
        mode = it % 4
        if mode == 0:
            b = b[:rnd.randrange(len(b))]
        elif mode == 1:
            for _ in range(rnd.randrange(1, 8)):
                b[rnd.randrange(len(b))] ^= 1 << rnd.randrange(8)
        elif mode == 2:
            i = rnd.randrange(len(b))
            b[i:i + 4] = b"\xff\xff\xff\xff"
        else:
            i = rnd.randrange(len(b))
            b = b[:i] + bytes(rnd.randrange(256) for _ in range(rnd.randrange(1, 64))) + b[i:]
        p2 = str(tmp_path / f"f{it}")
        open(p2 + ".index", "wb").write(bytes(b))
        shutil.copy(prefix + ".data-00000-of-00001", p2 + ".data-00000-of-00001")
        try:
            with bundle.BundleReader(p2) as r:
                opened += 1
                for k in r.keys():
                    try:
                        v = r.read(k)
                    except bundle.BundleError:
                        continue
                    if k in tensors and v.shape == tensors[k].shape:  # a tensor that passes its CRC is the original
                        assert np.array_equal(v, tensors[k]), k
        except bundle.BundleError:
            pass
    assert opened < 150  # the block checksums catch most of it
