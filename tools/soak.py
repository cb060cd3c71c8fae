#!/usr/bin/env python
"""Runs ON the GPU box: the same training run repeated many times through the resident loop; every repetition must end
with the same iteration count, b and alpha bits (the loop's exchanges are lock-free: a race would show up as a difference,
a lost message as a DPSVM_ERR_COMM after the bounded wait).  usage: soak.py [workload] [repetitions] [cache lines]"""
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dpsvm_b200 import capi, datagen  # noqa: E402


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "adult32k"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    cache = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    x, y, prm = datagen.make(wl)
    n, d = x.shape
    seen = set()
    with capi.Trainer(n, d, prm["c"], prm["gamma"], prm["epsilon"], max_iter=2000000, cache_size=cache) as t:
        t.upload(x, y)
        for r in range(reps):
            if r:
                t.reset()
            st = t.train()
            alpha, f, _ = t.download()
            seen.add((st.num_iter, st.converged, float(st.b), zlib.crc32(alpha.tobytes()), zlib.crc32(f.tobytes()), st.cache_hits))
    print("%s x %d (cache %d): %d distinct outcome(s): %s" % (wl, reps, cache, len(seen), sorted(seen)[:2]))
    assert len(seen) == 1


if __name__ == "__main__":
    main()
