"""MNIST input pipeline (C13) and its synthetic stand-in.

``input_data.read_data_sets(data_dir, one_hot=True)`` → ``mnist.train.next_batch(bs)`` returning
``(images[bs,784] float32 in [0,1], labels[bs,10] one-hot)`` (``encoder:24,154,164``,
``gan:29,177,187``, ``lstm:6,117,126-127,136-137``); 55 000 train / 5 000 validation / 10 000
test, reshuffled every epoch.  Reads the four idx(.gz) files from ``data_dir``; with
``synthetic=True`` (no dataset can be downloaded here) it generates MNIST-shaped data.
"""
from __future__ import annotations

import gzip
import os
import struct
from typing import Optional, Tuple

import numpy as np

FILES = {"train_images": "train-images-idx3-ubyte", "train_labels": "train-labels-idx1-ubyte",
         "test_images": "t10k-images-idx3-ubyte", "test_labels": "t10k-labels-idx1-ubyte"}
VALIDATION_SIZE = 5000


def _open(path: str):
    if os.path.exists(path + ".gz"):
        return gzip.open(path + ".gz", "rb")
    return open(path, "rb")


def read_idx_images(path: str) -> np.ndarray:
    with _open(path) as f:
        magic, n, rows, cols = struct.unpack(">IIII", f.read(16))
        if magic != 2051:
            raise ValueError("bad magic %d in %s" % (magic, path))
        data = np.frombuffer(f.read(n * rows * cols), dtype=np.uint8)
    return (data.reshape(n, rows * cols).astype(np.float32) / 255.0)


def read_idx_labels(path: str, one_hot: bool, num_classes: int = 10) -> np.ndarray:
    with _open(path) as f:
        magic, n = struct.unpack(">II", f.read(8))
        if magic != 2049:
            raise ValueError("bad magic %d in %s" % (magic, path))
        lab = np.frombuffer(f.read(n), dtype=np.uint8).astype(np.int64)
    if not one_hot:
        return lab
    out = np.zeros((n, num_classes), dtype=np.float32)
    out[np.arange(n), lab] = 1.0
    return out


class DataSet:
    """``mnist.train`` / ``.validation`` / ``.test``: ``images``, ``labels``, ``next_batch``."""

    def __init__(self, images: np.ndarray, labels: np.ndarray, seed: int = 0):
        assert images.shape[0] == labels.shape[0]
        self._images, self._labels = images, labels
        self._n = images.shape[0]
        self._index = 0
        self.epochs_completed = 0
        self._rng = np.random.RandomState(seed)
        self._perm = self._rng.permutation(self._n)

    @property
    def images(self) -> np.ndarray:
        return self._images

    @property
    def labels(self) -> np.ndarray:
        return self._labels

    @property
    def num_examples(self) -> int:
        return self._n

    def next_batch(self, batch_size: int) -> Tuple[np.ndarray, np.ndarray]:
        if self._index + batch_size > self._n:
            # finish the epoch with what is left, reshuffle, top up from the new epoch
            rest = self._perm[self._index:]
            self.epochs_completed += 1
            self._perm = self._rng.permutation(self._n)
            need = batch_size - rest.shape[0]
            idx = np.concatenate([rest, self._perm[:need]])
            self._index = need
        else:
            idx = self._perm[self._index:self._index + batch_size]
            self._index += batch_size
        return self._images[idx], self._labels[idx]


class Datasets:
    def __init__(self, train: DataSet, validation: DataSet, test: DataSet):
        self.train, self.validation, self.test = train, validation, test


def synthetic_mnist(n_train: int = 8192, n_test: int = 1024, seed: int = 0, one_hot: bool = True) -> Datasets:
    """MNIST-shaped data with class structure (a noisy per-class template), so that a
    classifier can actually learn from it."""
    rng = np.random.RandomState(seed)
    templates = (rng.rand(10, 784) > 0.7).astype(np.float32)

    def make(n):
        lab = rng.randint(0, 10, size=n)
        img = np.clip(templates[lab] * 0.8 + rng.rand(n, 784).astype(np.float32) * 0.3, 0.0, 1.0).astype(np.float32)
        if one_hot:
            oh = np.zeros((n, 10), dtype=np.float32)
            oh[np.arange(n), lab] = 1.0
            return img, oh
        return img, lab

    tr_i, tr_l = make(n_train)
    va_i, va_l = make(max(64, n_test // 2))
    te_i, te_l = make(n_test)
    return Datasets(DataSet(tr_i, tr_l, seed + 1), DataSet(va_i, va_l, seed + 2), DataSet(te_i, te_l, seed + 3))


def read_data_sets(data_dir: str, one_hot: bool = False, synthetic: bool = False, seed: int = 0,
                   validation_size: int = VALIDATION_SIZE) -> Datasets:
    if synthetic:
        return synthetic_mnist(seed=seed, one_hot=one_hot)
    paths = {k: os.path.join(data_dir, v) for k, v in FILES.items()}
    missing = [p for p in paths.values() if not (os.path.exists(p) or os.path.exists(p + ".gz"))]
    if missing:
        raise FileNotFoundError(
            "MNIST files not found in %s (%s). The reference bakes them into its image "
            "(encoder/Dockerfile:3); pass --synthetic to train on MNIST-shaped synthetic data."
            % (data_dir, ", ".join(os.path.basename(m) for m in missing)))
    tr_i = read_idx_images(paths["train_images"])
    tr_l = read_idx_labels(paths["train_labels"], one_hot)
    te_i = read_idx_images(paths["test_images"])
    te_l = read_idx_labels(paths["test_labels"], one_hot)
    v = validation_size
    return Datasets(DataSet(tr_i[v:], tr_l[v:], seed + 1), DataSet(tr_i[:v], tr_l[:v], seed + 2),
                    DataSet(te_i, te_l, seed + 3))
