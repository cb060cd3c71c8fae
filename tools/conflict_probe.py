#!/usr/bin/env python
"""Runs ON the GPU box: what the shared-memory bank conflicts of the sparse sweep cost.  Two 60 000 x 784 sets with the
same number of non-zeros per row, one with random columns and one whose half-warps read 16 different bank pairs at
every step by construction; prints us per iteration for each (profiles/r01_sparse_ncu_summary.md)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dpsvm_b200 import capi
n, d = 60000, 784
rng = np.random.default_rng(0)
y = np.where(rng.random(n) < 0.5, 1, -1).astype(np.int32)
vals = rng.random((n, d), dtype=np.float32) * 0.5 + 0.25
res = np.arange(d) % 16
blk = (np.arange(d) // 16) % 4 != 3
mask_struct = np.where((np.arange(n) % 2 == 0)[:, None], (res < 8)[None, :], (res >= 8)[None, :]) & blk[None, :]
mask_rand = rng.random((n, d)) < mask_struct.mean()
os.environ["DPSVM_LAYOUT"] = "sparse"
for name, m in (("structured (conflict-free half-warps)", mask_struct), ("random columns", mask_rand), ("structured again", mask_struct)):
    x = np.ascontiguousarray(vals * m, dtype=np.float32)
    t = capi.Trainer(n, d, 10.0, 0.02, 1e-3, max_iter=10**8, cache_size=0)
    t.upload(x, y)
    lanes, nnz, sweep = t.layout_info()
    t.train_async(512); t.sync()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    t.train_async(4096); st = t.sync()
    dt = time.perf_counter() - t0
    print("%-40s lanes %d nnz/row %.1f sweep %.1f MB: %.2f us/iteration (converged %d)" % (name, lanes, nnz / n, sweep / 1e6, 1e6 * dt / 4096, st.converged), flush=True)
    t.close()
