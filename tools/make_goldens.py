#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the REFERENCE's own seq.cpp (oracle/_ref/libseqref.so,
built from /root/reference by oracle/Makefile) on seeded synthetic inputs.

Run in the build container (needs /root/reference):   python tools/make_goldens.py
The GPU box has no /root/reference; tests there read these committed files.

Provenance of every vector: reference seq.cpp functions, g++ 13.3 -O3 -std=c++11, x86-64 SSE2, BLAS
through oracle/cblas_shim/cblas.h (netlib saxpy / scale-ssq snrm2 restated; the reference's BLAS is
not vendored).  Inputs are NOT stored: they are regenerated from (generator, n, d, seed) by
dpsvm_b200/datagen.py and checked against the stored SHA-256.
"""
import hashlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dpsvm_b200 import datagen  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")

PROVENANCE = ("reference seq.cpp via oracle/ref_harness.cpp; g++ -O3 -std=c++11; "
              "BLAS = oracle/cblas_shim (netlib algorithms restated)")


def dups_case(seed=7, n=96, d=16):
    """Exact duplicate rows with opposite labels -> eta == 0 iterations."""
    rng = np.random.default_rng([seed, 9])
    x = rng.random((n, d), dtype=np.float32)
    y = np.where(rng.random(n) < 0.5, 1, -1).astype(np.int32)
    x[16:32] = x[0:16]
    y[16:32] = -y[0:16]
    return x, y


CASES = {
    # name: (generator, gen kwargs, c, gamma, eps, max_iter, n_test)
    "mnist_512": ("mnist_like", dict(n=512, d=784, seed=11), 10.0, 0.25, 1e-3, 150000, 128),
    "adult_2048": ("adult_like", dict(n=2048, seed=1), 100.0, 0.5, 1e-3, 150000, 256),
    "mnist_4096": ("mnist_like", dict(n=4096, d=784, seed=1), 10.0, 0.25, 1e-3, 150000, 0),
    "mnist_1024_tight": ("mnist_like", dict(n=1024, d=784, seed=12), 10.0, 0.25, 1e-5, 150000, 128),
    "adult_1000_tight": ("adult_like", dict(n=1000, seed=3), 100.0, 0.5, 1e-5, 400000, 128),
    "odd_d_333x37": ("mnist_like", dict(n=333, d=37, seed=4), 10.0, 0.25, 1e-3, 150000, 0),
    "dups_96x16": ("dups", dict(), 10.0, 0.5, 1e-3, 20000, 0),
    "tiny_3x5": ("mnist_like", dict(n=3, d=5, seed=2), 1.0, 0.25, 1e-3, 1000, 0),
    # the 1 % iteration bound above 4 096 rows, at a tolerance that pins the optimum (minutes of seq.cpp on one core)
    "mnist_8192_tight": ("mnist_like", dict(n=8192, d=784, seed=13), 10.0, 0.25, 1e-5, 400000, 256),
}
TRACE_CAP = {"mnist_8192_tight": 2048}      # pairs kept in the file (the others keep every iteration)


def gen(name, kw):
    if name == "mnist_like":
        return datagen.mnist_like(**kw)
    if name == "adult_like":
        return datagen.adult_like(**kw)
    if name == "dups":
        return dups_case()
    raise ValueError(name)


def gen_test(name, kw, n_test):
    kw = dict(kw)
    kw["n"] = n_test
    kw["seed"] = kw.get("seed", 0) + 1000
    return gen(name, kw)


def sha(x, y):
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(x).tobytes())
    h.update(np.ascontiguousarray(y).tobytes())
    return h.hexdigest()


def main():
    os.makedirs(OUT, exist_ok=True)
    only = set(sys.argv[1:])
    for name, (g, kw, c, gamma, eps, max_iter, n_test) in CASES.items():
        if only and name not in only:
            continue
        x, y = gen(g, kw)
        t = time.time()
        r = po.ref_train(x, y, c, gamma, eps, max_iter, trace_cap=TRACE_CAP.get(name, max_iter))
        dt = time.time() - t
        out = dict(
            generator=g, gen_kwargs=repr(kw), c=np.float32(c), gamma=np.float32(gamma), epsilon=np.float32(eps),
            max_iter=np.int32(max_iter), input_sha256=sha(x, y), n=np.int32(x.shape[0]), d=np.int32(x.shape[1]),
            alpha=r.alpha, f=r.f, num_iter=np.int32(r.num_iter), b_hi=r.b_hi, b_lo=r.b_lo, b=r.b,
            trace_idx=r.trace_idx, trace_b=r.trace_b, rule_disagreements=np.int32(r.rule_disagreements),
            provenance=PROVENANCE)
        if n_test:
            xt, yt = gen_test(g, kw, n_test)
            dual, lab = po.ref_predict(x, y, r.alpha, gamma, xt)
            out.update(test_sha256=sha(xt, yt), test_dual=dual, test_labels=lab, n_test=np.int32(n_test))
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        print("%-18s n=%d d=%d iters=%d b=%.6f nsv=%d nan=%d disagree=%d  (%.1fs)" % (
            name, x.shape[0], x.shape[1], r.num_iter, r.b, int((r.alpha != 0).sum()),
            int(np.isnan(r.alpha).sum()), r.rule_disagreements, dt))

    # Known-answer vector the reference itself carries: testblas.c:21-48 / mpi_sample.cpp:5-40,
    # a 3x3 row-major gemv whose printed answer is (-1, 3, -3).
    np.savez(os.path.join(OUT, "kat_gemv3x3.npz"),
             m=np.array([[3, 1, 3], [1, 5, 9], [2, 6, 5]], np.float32),
             x=np.array([-1, -1, 1], np.float32), y=np.array([-1, 3, -3], np.float32),
             provenance="testblas.c:21-48 (data) ; answer = cblas_dgemv RowMajor NoTrans")


if __name__ == "__main__":
    main()
