"""MNIST from local files, with an ``.npz`` cache (the job of the reference's
chainer/mnist_dataset.py:8-38 and chainer/mnist_helper.py:9-53, and of
``datasets.mnist.load_data(path=...)`` in tensorflow2/mnist_single.py:36-38).

Reads the four idx ``.gz`` files (or an ``mnist.npz``) from a directory; the idx payload is
decoded with one ``numpy.frombuffer`` per file instead of a byte-at-a-time loop.  There is no
network on the build machines, so when nothing is found ``synthetic=True`` returns random
MNIST-shaped data instead of downloading.
"""
from __future__ import annotations

import gzip
import os
import struct

import numpy as np

FILES = {"train_x": "train-images-idx3-ubyte.gz", "train_y": "train-labels-idx1-ubyte.gz",
         "test_x": "t10k-images-idx3-ubyte.gz", "test_y": "t10k-labels-idx1-ubyte.gz"}


def read_idx(path):
    """Decodes one idx file (gzip or plain) into a uint8 array of its declared shape."""
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rb") as f:
        raw = f.read()
    zero, dtype_code, ndim = struct.unpack(">HBB", raw[:4])
    if zero != 0 or dtype_code != 0x08:
        raise ValueError(f"{path}: not an unsigned-byte idx file")
    shape = struct.unpack(">" + "I" * ndim, raw[4: 4 + 4 * ndim])
    return np.frombuffer(raw, dtype=np.uint8, offset=4 + 4 * ndim).reshape(shape)


def make_npz(directory, out="mnist.npz"):
    arrays = {k: read_idx(os.path.join(directory, v)) for k, v in FILES.items()}
    path = os.path.join(directory, out)
    np.savez_compressed(path, x_train=arrays["train_x"], y_train=arrays["train_y"],
                        x_test=arrays["test_x"], y_test=arrays["test_y"])
    return path


def synthetic_mnist(n_train=2048, n_test=512, seed=0):
    rng = np.random.default_rng(seed)
    mk = lambda n: (rng.integers(0, 256, (n, 28, 28), dtype=np.uint8), rng.integers(0, 10, n).astype(np.uint8))  # noqa: E731
    return mk(n_train), mk(n_test)


def load_mnist(directory, synthetic=True):
    """Returns ((x_train, y_train), (x_test, y_test)) as uint8 arrays [N,28,28] / [N]."""
    npz = os.path.join(directory, "mnist.npz")
    if not os.path.exists(npz) and all(os.path.exists(os.path.join(directory, v)) for v in FILES.values()):
        npz = make_npz(directory)
    if os.path.exists(npz):
        d = np.load(npz)
        return (d["x_train"], d["y_train"]), (d["x_test"], d["y_test"])
    if not synthetic:
        raise FileNotFoundError(f"no MNIST files under {directory}")
    return synthetic_mnist()


def get_mnist(directory="./data/mnist", ndim=1, scale=1.0, synthetic=True):
    """Chainer's ``get_mnist``: lists of (float32 image, int label); ``ndim`` 1 -> 784, 3 -> 1x28x28."""
    (xt, yt), (xv, yv) = load_mnist(directory, synthetic)

    def prep(x, y):
        x = x.astype(np.float32) * (scale / 255.0)
        x = x.reshape(len(x), -1) if ndim == 1 else x.reshape(len(x), 1, 28, 28)
        return list(zip(x, y.astype(np.int64)))

    return prep(xt, yt), prep(xv, yv)
