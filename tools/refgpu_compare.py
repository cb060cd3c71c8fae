#!/usr/bin/env python
"""Runs ON the GPU box: the reference's own GPU path (svmTrain.cu + cache.cu, cuBLAS gemv +
Thrust, rebuilt unmodified for sm_100a as oracle/_ref/svmTrain_refgpu) against this repo's CLI on the
same CSV.  Loop time of each program = time between its "SETUP DONE" and "TOTAL TIME TAKEN" stdout lines."""
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dpsvm_b200 import datagen  # noqa: E402


def timed_run(cmd):
    # line-buffer the child (cout is synchronised with stdio), otherwise all lines arrive at exit
    p = subprocess.Popen(["stdbuf", "-oL"] + cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, bufsize=1)
    t_setup = t_total = None
    iters = None
    for line in p.stdout:
        now = time.perf_counter()
        if line.startswith("SETUP DONE"):
            t_setup = now
        elif line.startswith("TOTAL TIME TAKEN"):
            t_total = now
        elif "iteration number" in line or "Could not converge in" in line:
            iters = int([w for w in line.replace(".", " ").split() if w.isdigit()][0])
    p.wait()
    return (t_total - t_setup) if (t_setup and t_total) else None, iters, p.returncode


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
    cap = int(sys.argv[2]) if len(sys.argv) > 2 else 4000
    x, y = datagen.mnist_like(n, 784, seed=0)
    import pandas as pd
    t0 = time.perf_counter()
    df = pd.DataFrame(x)
    df.insert(0, "y", y)
    csv = "/tmp/refgpu_train.csv"
    df.to_csv(csv, header=False, index=False, float_format="%.9g")
    print("csv written in %.1f s (%.0f MB)" % (time.perf_counter() - t0, os.path.getsize(csv) / 1e6), flush=True)
    common = ["-a", "784", "-x", str(n), "-f", csv, "-c", "10", "-g", "0.25", "-e", "0.001", "-n", str(cap), "-s", "10"]
    rows = []
    for name, exe, extra in (("reference GPU path (cuBLAS gemv + Thrust), unmodified, sm_100a", os.path.join(ROOT, "oracle/_ref/svmTrain_refgpu"), []),
                             ("this repo, unmodified reference main() on libdpsvm (host loop)", os.path.join(ROOT, "build/svmTrain_refmain"), []),
                             ("this repo, CLI --host-loop (train_step1/2 per iteration)", os.path.join(ROOT, "dpsvm_b200/svmTrain"), ["--host-loop"]),
                             ("this repo, CLI (device-resident loop)", os.path.join(ROOT, "dpsvm_b200/svmTrain"), [])):
        if not os.path.exists(exe):
            continue
        secs, iters, rc = timed_run([exe] + common + ["-m", "/tmp/refgpu_model.txt"] + extra)
        rows.append((name, iters, secs, rc))
        print(name, iters, secs, rc, flush=True)
    print("\n| program (same CSV, %d×784, c=10 γ=0.25, first %d iterations) | iterations | loop s | iterations/s |" % (n, cap))
    print("|---|---|---|---|")
    for name, iters, secs, rc in rows:
        print("| %s | %s | %s | %s |" % (name, iters, ("%.3f" % secs) if secs else "–", ("%.0f" % (iters / secs)) if secs and iters else "rc=%d" % rc))


if __name__ == "__main__":
    main()
