#!/usr/bin/env python
"""Runs ON the GPU box: time dpsvm_predict (train accuracy of a converged model, get_train_accuracy's job)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dpsvm_b200 import capi, datagen

for wl in ("adult32k", "mnist60k"):
    x, y, prm = datagen.make(wl)
    n, d = x.shape
    with capi.Trainer(n, d, prm["c"], prm["gamma"], prm["epsilon"], max_iter=2000000, cache_size=n) as t:
        t.upload(x, y)
        st = t.train()
        alpha, _, _ = t.download(f=False)
        t0 = time.perf_counter()
        dual, pred, nsv, acc = t.predict(alpha, y, st.b, True)
        dt = time.perf_counter() - t0
        flops = 2.0 * n * nsv * d
        print("%s: predict on %d train rows x %d SVs x %d attrs: %.3f s (%.2f TFLOP/s fp32), accuracy %.4f" % (wl, n, nsv, d, dt, flops / dt / 1e12, acc))
