"""Input pipeline: MNIST loading with the reference's shared-filesystem guards, a
synthetic stand-in for boxes with no network, and the infinite shuffled
minibatch generator.

Reference behaviour (``horovod/tensorflow_mnist.py:76-85,92-115``; duplicated at
``horovod/tensorflow_mnist_gpu.py:103-124``):

* ``~/.keras/datasets`` is created tolerating EEXIST — ranks share a filesystem;
* each rank loads into its OWN file name ``MNIST-data-<rank>`` so concurrent
  writers never collide;
* pixels are scaled to [0,1] and flattened to 784;
* every rank sees the FULL dataset and shuffles it independently (no sharding);
* the generator reshuffles each pass and drops the short tail batch.
"""
from __future__ import annotations

import errno
import os
from typing import Iterator, Optional, Tuple

import numpy as np

Arrays = Tuple[Tuple[np.ndarray, np.ndarray], Tuple[np.ndarray, np.ndarray]]


def ensure_cache_dir(cache_dir: Optional[str] = None) -> str:
    cache_dir = cache_dir or os.path.join(os.path.expanduser("~"), ".keras", "datasets")
    if not os.path.exists(cache_dir):
        try:
            os.makedirs(cache_dir)
        except OSError as e:                       # another rank won the race
            if e.errno == errno.EEXIST and os.path.isdir(cache_dir):
                pass
            else:
                raise
    return cache_dir


def synthetic_mnist(n_train: int = 60000, n_test: int = 10000, seed: int = 1234) -> Arrays:
    """Deterministic, learnable 28x28 uint8 digits-like data: each class is a fixed
    random blob pattern plus per-sample noise and a small shift.  Same shapes and
    dtypes as ``keras.datasets.mnist.load_data``."""
    rng = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:28, 0:28]
    templates = np.zeros((10, 28, 28), np.float32)
    for c in range(10):
        for _ in range(4):
            cy, cx = rng.uniform(6, 22, size=2)
            s = rng.uniform(2.0, 4.0)
            templates[c] += np.exp(-((yy - cy) ** 2 + (xx - cx) ** 2) / (2 * s * s))
        templates[c] /= templates[c].max()

    def make(n, r):
        y = r.randint(0, 10, size=n).astype(np.uint8)
        x = templates[y]
        sh = r.randint(-2, 3, size=(n, 2))
        out = np.empty((n, 28, 28), np.float32)
        for k in range(5):
            for l in range(5):
                sel = (sh[:, 0] == k - 2) & (sh[:, 1] == l - 2)
                if sel.any():
                    out[sel] = np.roll(x[sel], (k - 2, l - 2), axis=(1, 2))
        out += r.normal(0, 0.15, size=out.shape).astype(np.float32)
        return (np.clip(out, 0, 1) * 255).astype(np.uint8), y

    return make(n_train, rng), make(n_test, np.random.RandomState(seed + 1))


def load_data(path: str = "mnist.npz", *, cache_dir: Optional[str] = None, synthetic: Optional[bool] = None,
              n_train: int = 60000, n_test: int = 10000) -> Arrays:
    """``keras.datasets.mnist.load_data(path)`` equivalent.  ``path`` is a file name
    inside the cache dir (the reference passes ``'MNIST-data-%d' % rank``).  With no
    network, the file must already exist (or ``B200DP_MNIST_NPZ`` point at a Keras
    ``mnist.npz``); otherwise synthetic data is produced and — like a download —
    cached under the per-rank name."""
    cache = ensure_cache_dir(cache_dir)
    full = path if os.path.isabs(path) else os.path.join(cache, path)
    src = os.environ.get("B200DP_MNIST_NPZ")
    if synthetic is None:
        synthetic = os.environ.get("B200DP_SYNTHETIC", "") == "1"
    if not synthetic:
        for cand in (full, full + ".npz", src):
            if cand and os.path.exists(cand):
                with np.load(cand, allow_pickle=False) as f:
                    return (f["x_train"], f["y_train"]), (f["x_test"], f["y_test"])
    (xtr, ytr), (xte, yte) = synthetic_mnist(n_train, n_test)
    if not synthetic:
        tmp = f"{full}.tmp.{os.getpid()}.npz"
        np.savez(tmp, x_train=xtr, y_train=ytr, x_test=xte, y_test=yte)
        os.replace(tmp, full if full.endswith(".npz") else full + ".npz")
    return (xtr, ytr), (xte, yte)


def preprocess(x: np.ndarray) -> np.ndarray:
    """(-1,28,28) uint8 -> (-1,784) float32 in [0,1]."""
    return (np.reshape(x, (-1, 784)).astype(np.float32)) / np.float32(255.0)


def train_input_generator(x_train: np.ndarray, y_train: np.ndarray, batch_size: int = 64,
                          seed: Optional[int] = None) -> Iterator[Tuple[np.ndarray, np.ndarray]]:
    """Infinite generator; fresh permutation every pass; tail dropped.  Unseeded by
    default, so ranks draw different orders (the reference's behaviour).  With the native loader the yielded arrays
    alias a ring slot and stay valid until four more batches have been drawn (copy them to keep them longer)."""
    assert len(x_train) == len(y_train)
    rng = np.random.RandomState(seed) if seed is not None else np.random
    if os.environ.get("B200DP_NATIVE_LOADER", "1") != "0" and x_train.dtype == np.float32 and 1 <= batch_size <= len(x_train):
        # same stream semantics, rows gathered by the native producer thread while the caller trains on the previous batch
        try:
            from .loader import BatchLoader
            ld = BatchLoader(x_train, y_train, batch_size, seed=int(rng.randint(0, 2 ** 31 - 1)), nslots=8, hold=4)
        except (RuntimeError, OSError, ValueError):
            ld = None
        if ld is not None and ld.native:
            try:
                while True:
                    xb, yb = ld.next()
                    yield xb.numpy(), yb.numpy()            # views of a ring slot: valid until four more batches were drawn
            finally:
                ld.close()
    # pure-Python twin of the native stream: one index permutation per pass and a cursor into it; the arrays themselves are
    # never reordered, a batch is one fancy-index gather, and the rows that do not fill a last batch are skipped
    n_full = len(x_train) // batch_size
    while True:
        order = rng.permutation(len(x_train))
        for b in range(n_full):
            rows = order[b * batch_size:(b + 1) * batch_size]
            yield x_train[rows], y_train[rows]
