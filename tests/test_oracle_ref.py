"""Pins the oracle against the REFERENCE ITSELF (oracle/_ref, built from /root/reference/seq.cpp by
oracle/Makefile).  CPU only; skipped where the reference build is absent."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

REF_DIR = os.path.join(ROOT, "oracle", "_ref")
pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REF_DIR, "libseqref.so")),
                                reason="oracle/_ref not built (needs /root/reference)")


def test_port_equals_reference_functions(oracle):
    from dpsvm_b200 import datagen
    x, y = datagen.mnist_like(300, 96, seed=21)
    r = oracle.ref_train(x, y, 10.0, 0.25, 1e-3, trace_cap=100000)
    p = oracle.train(x, y, 10.0, 0.25, 1e-3, trace_cap=100000)
    assert r.num_iter == p.num_iter and r.num_iter > 50
    assert np.array_equal(r.trace_idx, p.trace_idx)
    assert np.array_equal(r.alpha, p.alpha) and np.array_equal(r.f, p.f)
    assert r.b == p.b


def test_rbf_kernel_matches_reference(oracle):
    rng = np.random.default_rng(0)
    for d in (1, 5, 123, 784):
        a = rng.random(d, dtype=np.float32)
        b = rng.random(d, dtype=np.float32)
        assert oracle.rbf(a, b, 0.25) == oracle.ref_rbf(a, b, 0.25)
        assert oracle.ref_rbf(a, a, 0.25) == 1.0


def test_predict_matches_reference(oracle, golden):
    g = golden("mnist_512")
    dual, lab = oracle.predict(g["x"], g["y"], g["alpha"], float(g["gamma"]), 0.0, g["xt"], convention=0)
    assert np.array_equal(dual, g["test_dual"]) and np.array_equal(lab, g["test_labels"])


@pytest.mark.skipif(not os.path.exists(os.path.join(REF_DIR, "seq")), reason="seq binary not built")
def test_unmodified_seq_binary_agrees(oracle, tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    """The byte-for-byte unmodified reference program (its own main, CSV loader and model writer)."""
    from dpsvm_b200 import datagen
    x, y = datagen.mnist_like(400, 64, seed=5)
    # seq.cpp keeps file names in char[60] (seq.cpp:38-39): run in tmp_path with relative names
    csv, model = "s.csv", "s.model"
    datagen.write_csv(csv, x, y)
    out = subprocess.run([os.path.join(REF_DIR, "seq"), "-a", "64", "-x", "400", "-f", csv, "-c", "10", "-g", "0.25",
                          "-e", "0.001", "-m", model], capture_output=True, text=True, check=True).stdout
    lines = [l for l in out.splitlines() if not l.startswith("Current iteration")]
    iters = int([l for l in lines if l.startswith("Converged at iteration number")][0].split(":")[1])
    b = float([l for l in lines if l.startswith("b:")][0].split(":")[1])
    p = oracle.train(x, y, 10.0, 0.25, 1e-3)
    assert iters == p.num_iter
    assert abs(b - float(p.b)) < 1e-5                 # printed with 6 significant digits
    model_lines = open(model).read().splitlines()
    assert float(model_lines[0]) == 0.25              # gamma line, then one line per alpha != 0 (seq.cpp:302-313)
    assert len(model_lines) - 1 == int((p.alpha != 0).sum())
    first_sv = int(np.flatnonzero(p.alpha != 0)[0])
    assert model_lines[1].split(",")[0] == ("%g" % p.alpha[first_sv])
