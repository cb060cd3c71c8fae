#!/usr/bin/env python
"""Runs ON the GPU box: times the fused pass kernel alone (stateless C-ABI launcher) over X sizes and
L2 strategies -- forward sweep, serpentine (alternate direction per launch), L2 eviction-priority
pinning of a prefix of X -- to locate the effective L2 capacity and pick defaults.
Writes JSON lines to gpurun_out/pass_sweep.jsonl."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dpsvm_b200 import capi  # noqa: E402


def time_pass(n, d, variants, reps=120, warm=12, lines=False):
    pitch = capi.pitch_for(d)
    x = torch.rand(n, pitch, device="cuda", dtype=torch.float32)
    if pitch != d:
        x[:, d:] = 0
    xsq = torch.empty(n, device="cuda", dtype=torch.float32)
    y = (torch.randint(0, 2, (n,), device="cuda", dtype=torch.int32) * 2 - 1).to(torch.int32)
    alpha = torch.zeros(n, device="cuda", dtype=torch.float32)
    f = torch.zeros(n, device="cuda", dtype=torch.float32)
    keys = torch.full((2,), -1, dtype=torch.int64, device="cuda")
    ln = torch.empty(2, n, device="cuda", dtype=torch.float32) if lines else None
    capi.row_norms(x, n, d, pitch, xsq, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()

    def one(i):
        st = torch.cuda.current_stream().cuda_stream
        r_hi, r_lo = (7 * i + 1) % n, (13 * i + 5) % n
        capi.smo_pass(stream=st, d_x=x, n=n, d=d, pitch=pitch, d_xsq=xsq, d_y=y, d_alpha=alpha, d_f=f,
                      d_x_hi=x.data_ptr() + 4 * pitch * r_hi, d_x_lo=x.data_ptr() + 4 * pitch * r_lo, xsq_hi=1.0, xsq_lo=1.0,
                      delta_hi=0.0, delta_lo=0.0, gamma=0.25, c=10.0, index_base=0, d_keys=keys,
                      d_line_hi=ln[0] if lines else None, d_line_lo=ln[1] if lines else None,
                      variant=variants[i % len(variants)])
    # replay from a CUDA graph so that host launch overhead (~13 us per ctypes call) cannot be the floor
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for i in range(2):
            one(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(reps):
            one(i)
    g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / reps
    abytes = 4.0 * n * d + 20.0 * n + (8.0 * n if lines else 0)
    return us, abytes / us / 1e3


def main():
    out = open(os.path.join(ROOT, "gpurun_out", "pass_sweep.jsonl"), "w")
    d = int(os.environ.get("SWEEP_D", "784"))
    sizes = [8000, 16000, 24000, 32000, 40000, 48000, 60000, 90000, 120000]
    if len(sys.argv) > 1:
        sizes = [int(v) for v in sys.argv[1:]]
    for n in sizes:
        mb = 4e-6 * n * d
        modes = {"fwd": [0], "serp": [0, 1]}
        for pin in (32, 48, 64, 80, 96, 112):
            if pin < mb:
                modes["pin%d" % pin] = [2 | (pin << 8)]
                modes["pin%d+serp" % pin] = [2 | (pin << 8), 3 | (pin << 8)]
        for name, variants in modes.items():
            us, gbs = time_pass(n, d, variants)
            rec = {"n": n, "d": d, "x_mb": round(mb, 1), "mode": name, "us": round(us, 2), "alg_gbs": round(gbs, 1)}
            print(json.dumps(rec), flush=True)
            out.write(json.dumps(rec) + "\n")
    out.close()


if __name__ == "__main__":
    main()
