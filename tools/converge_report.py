#!/usr/bin/env python
"""Runs ON the GPU box: time-to-converge of full training runs (BASELINE metric, second half) through
the C ABI, as a markdown table.  usage: converge_report.py [workload ...]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dpsvm_b200 import capi, datagen  # noqa: E402


def run(workload, cache, max_iter=2000000):
    x, y, prm = datagen.make(workload)
    n, d = x.shape
    t = capi.Trainer(n, d, prm["c"], prm["gamma"], prm["epsilon"], max_iter=max_iter, cache_size=cache)
    t0 = time.perf_counter()
    t.upload(x, y)
    t1 = time.perf_counter()
    st = t.train()
    alpha, _, _ = t.download(f=False)
    t2 = time.perf_counter()
    resident, chunk = t.loop_info()
    nsv = int((alpha > 0).sum())
    bound = int((alpha == np.float32(prm["c"])).sum())
    acc = None
    if n <= 40000:
        _, _, _, acc = t.predict(alpha, y, st.b, subtract_b=True)
    t.close()
    hit = (100.0 * st.cache_hits / st.cache_lookups) if st.cache_lookups else 0.0
    print("| %s | %d×%d | c=%g γ=%g ε=%g | %d | %s | %d | %s | %.3f | %.3f | %.0f | %.1f %% | %d (%d at C) | %s | %.6f |" % (
        workload, n, d, prm["c"], prm["gamma"], prm["epsilon"], cache, "resident" if resident else "graph", st.num_iter,
        "yes" if st.converged else "NO", t1 - t0, t2 - t1, st.num_iter / (t2 - t1), hit, nsv, bound,
        ("%.4f" % acc) if acc is not None else "–", st.b))
    sys.stdout.flush()


def main():
    wl = sys.argv[1:] or ["adult32k", "mnist60k"]
    print("| workload | shape | params | cache lines | loop | iterations | converged | upload+setup s | train s | iter/s | cache hit rate | #SV | train accuracy | b |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    for w in wl:
        n = {"adult32k": 32561, "mnist60k": 60000}.get(w, 0)
        for cache in (10, 0, n):
            if cache == n and n == 0:
                continue
            run(w, cache)


if __name__ == "__main__":
    main()
