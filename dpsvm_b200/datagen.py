"""Seeded synthetic training sets in the layouts the reference's converters produce.

The reference ships no data (``data/train.csv`` is absent from the checkout), so every
configuration in BASELINE.json runs on synthetic inputs of the same shape and value range:

* ``mnist_like``  -- ``scripts/convert_mnist_to_odd_even.py:24-26``: 784 pixels scaled to [0,1], label +1
  for an even class, -1 for an odd one.  SURVEY's recipe (prototypes 19 % non-zero plus clipped Gaussian
  noise on half the pixels) comes out at 38.7 % non-zero cells; real MNIST has 19 %.
* ``adult_like``  -- ``scripts/convert_adult.py:24-31``: 123 binary columns, column 0 always 0,
  14 one-hot groups over columns 1..122 (exactly 14 ones per row), about 24 % positives.
* ``blobs``       -- dense Gaussian mixture for the large sharded configurations.

Rows are generated in fixed blocks of ``BLOCK`` rows, each from its own ``SeedSequence`` child, so
any row range ``[row0, row0+n)`` can be produced on its own (one shard per GPU rank) and the data
does not depend on how many ranks there are.
"""
from __future__ import annotations

import numpy as np

BLOCK = 8192
ADULT_GROUPS = (5, 8, 5, 16, 5, 7, 14, 6, 5, 2, 5, 5, 5, 34)  # sums to 122


def _blocks(row0: int, n: int):
    b = row0 // BLOCK
    while b * BLOCK < row0 + n:
        lo = max(row0, b * BLOCK)
        hi = min(row0 + n, (b + 1) * BLOCK)
        yield b, lo - b * BLOCK, hi - b * BLOCK
        b += 1


def mnist_like(n: int, d: int = 784, seed: int = 0, row0: int = 0):
    """Returns (x float32 [n,d] C-contiguous, y int32 [n] in {+1,-1})."""
    prng = np.random.default_rng([seed, 0])
    protos = ((prng.random((20, d)) < 0.19) * prng.random((20, d))).astype(np.float32)
    x = np.empty((n, d), dtype=np.float32)
    y = np.empty(n, dtype=np.int32)
    out = 0
    for b, lo, hi in _blocks(row0, n):
        rng = np.random.default_rng([seed, 1, b])
        pid = rng.integers(0, 20, BLOCK)
        noise = rng.standard_normal((BLOCK, d), dtype=np.float32)
        mask = rng.random((BLOCK, d), dtype=np.float32) < 0.5
        flip = rng.random(BLOCK) < 0.01
        xb = protos[pid[lo:hi]] + np.float32(0.12) * noise[lo:hi] * mask[lo:hi]
        np.clip(xb, 0.0, 1.0, out=xb)
        yb = np.where(pid[lo:hi] % 2 == 0, 1, -1).astype(np.int32)
        yb[flip[lo:hi]] *= -1
        x[out:out + hi - lo] = xb
        y[out:out + hi - lo] = yb
        out += hi - lo
    return x, y


def adult_like(n: int, seed: int = 0):
    """Returns (x float32 [n,123] of 0/1, y int32 [n]).  Generated whole (the label threshold is a
    quantile of the whole set)."""
    rng = np.random.default_rng([seed, 2])
    d = 123
    w = rng.standard_normal(d)
    x = np.zeros((n, d), dtype=np.float32)
    col = 1
    for g in ADULT_GROUPS:
        p = rng.dirichlet(np.full(g, 0.7))
        cat = rng.choice(g, size=n, p=p)
        x[np.arange(n), col + cat] = 1.0
        col += g
    score = x.astype(np.float64) @ w + 0.8 * rng.standard_normal(n)
    y = np.where(score > np.quantile(score, 0.76), 1, -1).astype(np.int32)
    return x, y


def blobs(n: int, d: int = 128, seed: int = 3, row0: int = 0, clusters: int = 32, spread: float = 1.4):
    """Gaussian mixture: `clusters` centres at distance ~spread*sqrt(d)... scaled so that
    gamma = 1/d gives kernel arguments of order 1.  Label = parity of the cluster id, 2 % flipped."""
    prng = np.random.default_rng([seed, 0])
    centres = (spread * prng.standard_normal((clusters, d))).astype(np.float32)
    x = np.empty((n, d), dtype=np.float32)
    y = np.empty(n, dtype=np.int32)
    out = 0
    for b, lo, hi in _blocks(row0, n):
        rng = np.random.default_rng([seed, 1, b])
        cid = rng.integers(0, clusters, BLOCK)
        noise = rng.standard_normal((BLOCK, d), dtype=np.float32)
        flip = rng.random(BLOCK) < 0.02
        xb = centres[cid[lo:hi]] + noise[lo:hi]
        yb = np.where(cid[lo:hi] % 2 == 0, 1, -1).astype(np.int32)
        yb[flip[lo:hi]] *= -1
        x[out:out + hi - lo] = xb
        y[out:out + hi - lo] = yb
        out += hi - lo
    return x, y


def write_csv(path: str, x: np.ndarray, y: np.ndarray) -> None:
    """`label,f0,...,f{d-1}` per line -- the layout parse.cpp:10-43 reads; values round-trip fp32 bit for bit
    through stof().  Written by the host library's multi-threaded writer (dpsvm_host_write_csv)."""
    import ctypes
    import os
    lib = ctypes.CDLL(os.path.join(os.path.dirname(os.path.abspath(__file__)), "libdpsvm_csv.so"))
    lib.dpsvm_host_write_csv.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    x = np.ascontiguousarray(x, np.float32)
    y = np.ascontiguousarray(y, np.int32)
    if lib.dpsvm_host_write_csv(path.encode(), x.ctypes.data, y.ctypes.data, x.shape[0], x.shape[1]) != 0:
        raise OSError("could not write " + path)


def make(workload: str, n: int | None = None, seed: int | None = None, row0: int = 0):
    """Named workloads of BASELINE.json:configs. Returns (x, y, params dict)."""
    if workload == "mnist60k":
        x, y = mnist_like(n or 60000, 784, 0 if seed is None else seed, row0)
        return x, y, dict(c=10.0, gamma=0.25, epsilon=1e-3)
    if workload == "adult32k":
        x, y = adult_like(n or 32561, 0 if seed is None else seed)
        return x, y, dict(c=100.0, gamma=0.5, epsilon=1e-3)
    if workload == "mnist1m":
        x, y = mnist_like(n or 1000000, 784, 2 if seed is None else seed, row0)
        return x, y, dict(c=10.0, gamma=0.25, epsilon=1e-3)
    if workload == "blobs4m":
        x, y = blobs(n or 4000000, 128, 3 if seed is None else seed, row0)
        return x, y, dict(c=10.0, gamma=1.0 / 128.0, epsilon=1e-3)
    raise ValueError("unknown workload %r" % workload)
