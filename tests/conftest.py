import hashlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def _sha(x, y):
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(x).tobytes())
    h.update(np.ascontiguousarray(y).tobytes())
    return h.hexdigest()


def load_golden(name):
    """Golden outputs of the reference's own seq.cpp (tools/make_goldens.py) plus the regenerated
    inputs, verified against the stored SHA-256 so a drifting generator cannot go unnoticed."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import make_goldens as mg
    g = dict(np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False))
    gen, kw = str(g["generator"]), eval(str(g["gen_kwargs"]))
    x, y = mg.gen(gen, kw)
    assert _sha(x, y) == str(g["input_sha256"]), "synthetic generator no longer reproduces the golden inputs"
    g["x"], g["y"] = x, y
    if "n_test" in g:
        xt, yt = mg.gen_test(gen, kw, int(g["n_test"]))
        assert _sha(xt, yt) == str(g["test_sha256"])
        g["xt"], g["yt"] = xt, yt
    return g


@pytest.fixture(scope="session")
def golden():
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = load_golden(name)
        return cache[name]
    return get


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    pyoracle.port()
    return pyoracle
