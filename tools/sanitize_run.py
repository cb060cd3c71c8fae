#!/usr/bin/env python
"""Small workload for compute-sanitizer (one tool per gpurun call): both loops, cache on, predict,
the stateless launchers and the allocation-free entry point, on a 512x784 golden set."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_golden  # noqa: E402
from dpsvm_b200 import capi  # noqa: E402

g = load_golden("mnist_512")
x, y = np.ascontiguousarray(g["x"]), np.ascontiguousarray(g["y"])
for flags, cache in ((0, 10), (capi.FLAG_GRAPH_LOOP, 10), (capi.FLAG_GRAPH_LOOP, 512), (capi.FLAG_NO_GRAPH, 0)):
    with capi.Trainer(512, 784, 10.0, 0.25, 1e-3, cache_size=cache, flags=flags) as t:
        t.upload(x, y)
        st = t.train(120)
        rv, _ = t.step1()
        i_hi, i_lo = int(rv[0]), int(rv[1])
        a, f, _ = t.download()
        t.step2(i_hi, i_lo, float(a[i_hi]), float(a[i_lo]))
        dual, pred, nsv, acc = t.predict(a, y, st.b, True, t=g["xt"])
        print("flags", flags, "cache", cache, "iters", st.num_iter, "nsv", nsv, "hits", st.cache_hits)
print("sanitize workload done")
