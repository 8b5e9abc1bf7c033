"""Input pipeline: MNIST-shaped data sets, ``next_batch`` and the multi-thread queue reader.

The reference loads MNIST with ``input_data.read_data_sets`` (distribute_training.py:184) and feeds
``mnist.train.next_batch(128)`` through placeholders every step (:191-192, :224-226); its README mentions a
"ReadByQueue" variant that fills a queue from several reader threads (README.md:5-7). MNIST itself is not
available offline, so the default data set is synthetic with the same shapes and dtypes (images float32 in
[0,1] flattened to 784, labels int64, 55000 training examples) and enough class structure to be learnable,
which is what lets tests assert that the loss goes down.
"""
from __future__ import annotations

import gzip
import os
import queue
import struct
import threading
from typing import Iterator, Optional, Tuple

import numpy as np


class DataSet:
    """``mnist.train``-like object: shuffles each epoch, rolls over by concatenating the tail of one epoch with
    the head of the next (the TF tutorial loader's behaviour). Unlike the reference it is seeded."""

    def __init__(self, images: np.ndarray, labels: np.ndarray, seed: int = 0):
        assert images.shape[0] == labels.shape[0]
        self._images = np.ascontiguousarray(images, dtype=np.float32)
        self._labels = np.ascontiguousarray(labels, dtype=np.int64)
        self._rng = np.random.default_rng(seed)
        self._perm = np.arange(self.num_examples)
        self._index = 0
        self.epochs_completed = 0

    @property
    def num_examples(self) -> int:
        return self._images.shape[0]

    @property
    def images(self) -> np.ndarray:
        return self._images

    @property
    def labels(self) -> np.ndarray:
        return self._labels

    def epoch_permutation(self) -> np.ndarray:
        """The next epoch's example order, drawn from the data set's own generator — the sequence ``next_batch``
        consumes one permutation at a time; the native reader (:class:`NativePrefetchQueue`) takes it whole."""
        return self._rng.permutation(self.num_examples)

    def next_batch(self, batch_size: int, shuffle: bool = True) -> Tuple[np.ndarray, np.ndarray]:
        n = self.num_examples
        if self.epochs_completed == 0 and self._index == 0 and shuffle:
            self._perm = self._rng.permutation(n)
        start = self._index
        if start + batch_size > n:
            self.epochs_completed += 1
            rest = self._perm[start:]
            if shuffle:
                self._perm = self._rng.permutation(n)
            take = batch_size - rest.shape[0]
            self._index = take
            idx = np.concatenate([rest, self._perm[:take]])
        else:
            self._index = start + batch_size
            idx = self._perm[start:self._index]
        return self._images[idx], self._labels[idx]


def synthetic_mnist(num_examples: int = 55000, input_dim: int = 784, num_classes: int = 10, seed: int = 0,
                    noise: float = 0.35, shuffle_seed: Optional[int] = None) -> DataSet:
    """Class-conditional synthetic images: each class is a fixed sparse stroke pattern, each example that pattern
    at random intensity plus uniform noise, clipped to [0, 1]."""
    rng = np.random.default_rng(seed)
    templates = (rng.random((num_classes, input_dim)) < 0.2).astype(np.float32)
    labels = rng.integers(0, num_classes, size=num_examples, dtype=np.int64)
    gain = rng.uniform(0.5, 1.0, size=(num_examples, 1)).astype(np.float32)
    images = templates[labels] * gain + noise * rng.random((num_examples, input_dim), dtype=np.float32)
    np.clip(images, 0.0, 1.0, out=images)
    return DataSet(images, labels, seed=seed + 1 if shuffle_seed is None else shuffle_seed)


def _read_idx(path: str) -> np.ndarray:
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rb") as f:
        magic = struct.unpack(">I", f.read(4))[0]
        ndim = magic & 0xFF
        dims = struct.unpack(">" + "I" * ndim, f.read(4 * ndim))
        return np.frombuffer(f.read(), dtype=np.uint8).reshape(dims)


def load_mnist_idx(data_path: str, seed: int = 0, validation_size: int = 5000) -> DataSet:  # seed: shuffle order
    """Real MNIST from idx files, split like ``read_data_sets`` (55000 train after 5000 validation)."""
    def find(stem: str) -> str:
        for ext in ("", ".gz"):
            p = os.path.join(data_path, stem + ext)
            if os.path.exists(p):
                return p
        raise FileNotFoundError(os.path.join(data_path, stem))
    images = _read_idx(find("train-images-idx3-ubyte")).reshape(-1, 784).astype(np.float32) / 255.0
    labels = _read_idx(find("train-labels-idx1-ubyte")).astype(np.int64)
    return DataSet(images[validation_size:], labels[validation_size:], seed=seed)


def make_dataset(kind: str, data_path: str, num_examples: int, input_dim: int, num_classes: int, seed: int,
                 shuffle_seed: Optional[int] = None) -> DataSet:
    """``seed`` fixes the CONTENT of a synthetic set (every worker must see the same one, as with MNIST, :184);
    ``shuffle_seed`` the order in which this worker walks it."""
    if kind == "idx":
        return load_mnist_idx(data_path, seed if shuffle_seed is None else shuffle_seed)
    if kind == "synthetic":
        return synthetic_mnist(num_examples, input_dim, num_classes, seed, shuffle_seed=shuffle_seed)
    raise ValueError(f"unknown data kind {kind!r}")


class PrefetchQueue:
    """The README's multi-thread queue reader: ``threads`` producers call ``next_batch`` under a lock and park
    ready batches (optionally in pinned memory, ready for an asynchronous H2D copy) in a bounded queue."""

    def __init__(self, dataset: DataSet, batch_size: int, threads: int = 2, capacity: int = 8, pin: bool = False):
        self._ds, self._bs = dataset, batch_size
        self._q: "queue.Queue" = queue.Queue(maxsize=capacity)
        self._lock = threading.Lock()
        self._stop = threading.Event()
        self._pin = pin
        self._threads = [threading.Thread(target=self._work, daemon=True) for _ in range(max(1, threads))]
        for t in self._threads:
            t.start()

    def _work(self) -> None:
        import torch
        while not self._stop.is_set():
            with self._lock:
                xs, ys = self._ds.next_batch(self._bs)
            tx, ty = torch.from_numpy(np.ascontiguousarray(xs)), torch.from_numpy(np.ascontiguousarray(ys))
            if self._pin and torch.cuda.is_available():
                tx, ty = tx.pin_memory(), ty.pin_memory()
            while not self._stop.is_set():
                try:
                    self._q.put((tx, ty), timeout=0.1)
                    break
                except queue.Full:
                    continue

    def get(self, timeout: Optional[float] = 30.0):
        return self._q.get(timeout=timeout)

    def __iter__(self) -> Iterator:
        while True:
            yield self.get()

    def close(self) -> None:
        self._stop.set()
        for t in self._threads:
            t.join(timeout=2.0)


class NativePrefetchQueue:
    """Multi-thread reader in C++ (``csrc/io/loader.cc``): worker threads assemble batches into a ring of pinned
    host buffers without the GIL; ``get()`` costs a mutex hand-off instead of a 75 us numpy gather. The example
    order is the data set's own (``epoch_permutation``), so the batches equal ``DataSet.next_batch``'s.

    A returned batch stays valid for ``release_lag - 1`` further ``get()`` calls: long enough for the trainer's
    asynchronous H2D copy (one step ahead, plus ``--fetch_lag``) when the caller passes ``fetch_lag + 3``."""

    def __init__(self, dataset: DataSet, batch_size: int, threads: int = 2, release_lag: int = 3, pin: bool = True):
        import ctypes as C

        import torch

        from .io.bundle import lib
        self._ds, self._bs, self._lib, self._C = dataset, batch_size, lib(), C
        self._lag = max(1, release_lag)
        self._ring = self._lag + max(1, threads) + 2
        feat = dataset.images.shape[1]
        self._x = torch.empty(self._ring, batch_size, feat, dtype=torch.float32)
        self._y = torch.empty(self._ring, batch_size, dtype=torch.int64)
        if pin and torch.cuda.is_available():
            self._x, self._y = self._x.pin_memory(), self._y.pin_memory()
        xs = (C.c_void_p * self._ring)(*[self._x[i].data_ptr() for i in range(self._ring)])
        ys = (C.c_void_p * self._ring)(*[self._y[i].data_ptr() for i in range(self._ring)])
        self._h = self._lib.ttdio_loader_create(dataset.images.ctypes.data, dataset.labels.ctypes.data,
                                                dataset.num_examples, feat, batch_size, self._ring, max(1, threads), xs, ys)
        if not self._h:
            raise RuntimeError("ttdio_loader_create failed: " + self._lib.ttdio_last_error().decode(errors="replace"))
        self._taken = 0
        self._topup_every = max(1, min(64, dataset.num_examples // (2 * batch_size)))
        self._top_up()

    def _top_up(self) -> None:
        # keep at least one full epoch (and the ring's worth of batches) of indices ahead of the consumer
        need = max(self._ds.num_examples, (self._ring + 1 + self._topup_every) * self._bs)
        while self._lib.ttdio_loader_backlog(self._h) < need:
            perm = np.ascontiguousarray(self._ds.epoch_permutation(), dtype=np.int64)
            if self._lib.ttdio_loader_push_epoch(self._h, perm.ctypes.data, perm.shape[0]) != 0:
                raise RuntimeError("ttdio_loader_push_epoch failed: " + self._lib.ttdio_last_error().decode(errors="replace"))

    def get(self, timeout: Optional[float] = 30.0):
        """Next batch as (x, y) torch views of the ring slot (pinned on a GPU box)."""
        slot = self._lib.ttdio_loader_next(self._h, int((timeout or 30.0) * 1000))
        if slot < 0:
            raise queue.Empty() if slot == -1 else RuntimeError("native reader is closed")
        k = self._taken
        self._taken += 1
        if k - self._lag >= 0:
            self._lib.ttdio_loader_release(self._h, k - self._lag)
        if k % self._topup_every == 0:
            self._top_up()
        return self._x[slot], self._y[slot]

    def __iter__(self) -> Iterator:
        while True:
            yield self.get()

    def close(self) -> None:
        if self._h:
            self._lib.ttdio_loader_destroy(self._h)
            self._h = None
