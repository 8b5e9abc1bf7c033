"""Input pipeline: next_batch roll-over, the Python queue reader and the native (C++) multi-thread reader."""
import numpy as np
import pytest


def test_next_batch_rolls_over_like_the_tf_loader():
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    ds = synthetic_mnist(300, seed=4)
    seen = []
    for _ in range(7):  # 7 x 128 = 896 = almost three epochs of 300
        x, y = ds.next_batch(128)
        assert x.shape == (128, 784) and x.dtype == np.float32 and y.dtype == np.int64
        seen.append(y)
    assert ds.epochs_completed == 2
    first_epoch = np.concatenate(seen)[:300]
    assert np.array_equal(np.sort(first_epoch), np.sort(ds.labels))  # one epoch visits every example exactly once


@pytest.mark.parametrize("n,threads", [(1000, 1), (300, 3), (4096, 4)])
def test_native_reader_yields_exactly_next_batch(n, threads):
    """csrc/io/loader.cc: same example order as DataSet.next_batch (incl. the epoch roll-over), batches in order."""
    from tensorflow_train_distributed_b200.data import NativePrefetchQueue, synthetic_mnist
    a, b = synthetic_mnist(n, seed=3), synthetic_mnist(n, seed=3)
    q = NativePrefetchQueue(b, 128, threads=threads, release_lag=2)
    try:
        held = []
        for i in range(150):
            xa, ya = a.next_batch(128)
            xb, yb = q.get()
            assert np.array_equal(xa, xb.numpy()) and np.array_equal(ya, yb.numpy()), i
            held.append((xa, xb, i))
            if len(held) > 1:  # a batch stays intact for release_lag - 1 further get() calls
                xa0, xb0, j = held.pop(0)
                assert np.array_equal(xa0, xb0.numpy()), j
    finally:
        q.close()


def test_native_reader_rejects_bad_indices_and_closes_cleanly():
    import ctypes as C
    from tensorflow_train_distributed_b200.data import NativePrefetchQueue, synthetic_mnist
    from tensorflow_train_distributed_b200.io.bundle import lib
    q = NativePrefetchQueue(synthetic_mnist(256, seed=1), 64, threads=2)
    bad = np.array([0, 1, 256], dtype=np.int64)
    assert lib().ttdio_loader_push_epoch(q._h, bad.ctypes.data, 3) != 0
    assert b"out of range" in lib().ttdio_last_error()
    q.get()
    q.close()
    q.close()


def test_python_queue_reader_delivers_batches():
    from tensorflow_train_distributed_b200.data import PrefetchQueue, synthetic_mnist
    q = PrefetchQueue(synthetic_mnist(512, seed=2), 128, threads=2)
    try:
        for _ in range(6):
            x, y = q.get()
            assert tuple(x.shape) == (128, 784) and tuple(y.shape) == (128,)
    finally:
        q.close()


@pytest.mark.parametrize("gz", [False, True])
def test_mnist_idx_files_load_like_read_data_sets(tmp_path, gz):
    """``--data idx``: the loader of real MNIST files (distribute_training.py:184) — idx3/idx1 big-endian headers,
    plain or gzipped, images scaled to [0,1] and flattened, the first 5000 examples held out for validation."""
    import gzip
    import struct
    from tensorflow_train_distributed_b200.data import load_mnist_idx, make_dataset
    rng = np.random.default_rng(0)
    n = 5200
    imgs = rng.integers(0, 256, size=(n, 28, 28), dtype=np.uint8)
    labs = rng.integers(0, 10, size=n, dtype=np.uint8)
    opener = gzip.open if gz else open
    ext = ".gz" if gz else ""
    with opener(tmp_path / ("train-images-idx3-ubyte" + ext), "wb") as f:
        f.write(struct.pack(">IIII", 0x00000803, n, 28, 28) + imgs.tobytes())
    with opener(tmp_path / ("train-labels-idx1-ubyte" + ext), "wb") as f:
        f.write(struct.pack(">II", 0x00000801, n) + labs.tobytes())
    ds = load_mnist_idx(str(tmp_path), seed=1)
    assert ds.num_examples == 200 and ds.images.shape == (200, 784) and ds.images.dtype == np.float32
    assert np.allclose(ds.images[0], imgs[5000].reshape(-1) / 255.0) and ds.labels[3] == labs[5003]
    assert 0.0 <= ds.images.min() and ds.images.max() <= 1.0 and ds.labels.dtype == np.int64
    x, y = make_dataset("idx", str(tmp_path), 0, 784, 10, seed=1).next_batch(64)
    assert x.shape == (64, 784) and y.shape == (64,)
