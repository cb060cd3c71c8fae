#!/usr/bin/env python
"""Runs ON the GPU box: trains every golden case through the C ABI and prints the parity numbers
against the reference's seq.cpp goldens as a markdown table (-> profiles/parity_rNN.md)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_golden  # noqa: E402
from dpsvm_b200 import capi  # noqa: E402
from oracle import pyoracle  # noqa: E402

NAMES = ["mnist_512", "adult_2048", "mnist_4096", "odd_d_333x37", "mnist_1024_tight", "adult_1000_tight",
         "mnist_8192_tight", "dups_96x16", "tiny_3x5"]


def main():
    layout = os.environ.get("DPSVM_LAYOUT", "auto")
    print("layout of the first table: %s\n" % layout)
    print("| case | n×d | eps | iters ours / seq | Δiter % | \\|Δb\\| | max\\|Δα\\| | #SV ours / seq | held-out label mismatches | "
          "same pair as seq (leading iters) | same pair as port form-1 (leading iters) |")
    print("|---|---|---|---|---|---|---|---|---|---|---|")
    spread = []
    for name in NAMES:
        g = load_golden(name)
        n, d = int(g["n"]), int(g["d"])
        c, gamma, eps = float(g["c"]), float(g["gamma"]), float(g["epsilon"])
        with capi.Trainer(n, d, c, gamma, eps, max_iter=int(g["max_iter"]), cache_size=0) as t:
            t.upload(np.ascontiguousarray(g["x"]), np.ascontiguousarray(g["y"]))
            # pair trace, one iteration at a time
            trace = []
            while True:
                st = t.train(1)
                trace.append((st.i_hi, st.i_lo))
                if st.converged or st.num_iter >= int(g["max_iter"]):
                    break
            alpha, f, _ = t.download()
            mism = "-"
            if "xt" in g:
                dual, pred, nsv, _ = t.predict(alpha, g["y"], st.b, subtract_b=False, t=g["xt"])
                mism = str(int((pred != g["test_labels"]).sum())) + "/" + str(len(pred))
        ref_trace = [tuple(r) for r in g["trace_idx"]]
        lead = 0
        for a, b in zip(trace, ref_trace):
            if a != b:
                break
            lead += 1
        lead1 = "-"
        if n * d <= 1100000:
            r = pyoracle.train(g["x"], g["y"], c, gamma, eps, int(g["max_iter"]), kernel_form=1, select_rule=1,
                               nthreads=pyoracle.max_threads(), trace_cap=len(trace) + 10)
            l1 = 0
            for a, b in zip(trace, [tuple(q) for q in r.trace_idx]):
                if a != b:
                    break
                l1 += 1
            lead1 = "%d (port: %d iters)" % (l1, r.num_iter)
        ri = int(g["num_iter"])
        print("| %s | %d×%d | %g | %d / %d | %.2f | %.2e | %.2e | %d / %d | %s | %d | %s |" % (
            name, n, d, eps, st.num_iter, ri, 100.0 * abs(st.num_iter - ri) / ri, abs(st.b - float(g["b"])),
            float(np.abs(alpha - g["alpha"]).max()), int((alpha != 0).sum()), int((g["alpha"] != 0).sum()), mism, lead, lead1))
        sys.stdout.flush()
        # the two summation orders side by side (dense rows / sparse stream of the same rows)
        its = {}
        for mode in ("dense", "sparse"):
            os.environ["DPSVM_LAYOUT"] = mode
            with capi.Trainer(n, d, c, gamma, eps, max_iter=int(g["max_iter"]), cache_size=0) as t:
                t.upload(np.ascontiguousarray(g["x"]), np.ascontiguousarray(g["y"]))
                s2 = t.train()
                a2, _, _ = t.download()
                its[mode] = (s2.num_iter, float(s2.b), int((a2 != 0).sum()))
        os.environ["DPSVM_LAYOUT"] = layout
        spread.append((name, ri, its))
    print("\n| case | iterations seq.cpp | dense order | Δ % | sparse order | Δ % | dense vs sparse Δ % | \\|b_dense − b_sparse\\| | #SV dense / sparse |")
    print("|---|---|---|---|---|---|---|---|---|")
    for name, ri, its in spread:
        dn, sp = its["dense"], its["sparse"]
        print("| %s | %d | %d | %.2f | %d | %.2f | %.2f | %.1e | %d / %d |" % (
            name, ri, dn[0], 100.0 * abs(dn[0] - ri) / ri, sp[0], 100.0 * abs(sp[0] - ri) / ri,
            100.0 * abs(dn[0] - sp[0]) / ri, abs(dn[1] - sp[1]), dn[2], sp[2]))


if __name__ == "__main__":
    main()
