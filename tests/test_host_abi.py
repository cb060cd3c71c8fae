"""The C-ABI library without a GPU: it loads, exports every symbol include/dpsvm_cuda.h declares,
its host-only helpers agree with the oracle, and compute entry points fail loudly with no device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, _have_gpu
from dpsvm_b200 import capi


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "dpsvm_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dpsvm_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = capi.lib()
    declared = _declared_symbols()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(L, name), "libdpsvm.so does not export " + name
    assert sorted(capi.EXPORTS) == declared
    assert L.dpsvm_abi_version() == 1


def test_binding_constants_match_the_header():
    """Flags and status codes of the ctypes binding are the header's, value for value."""
    text = open(os.path.join(ROOT, "include", "dpsvm_cuda.h")).read()
    defs = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define\s+DPSVM_(FLAG_[A-Z_]+)\s+(-?\d+)", text)}
    assert set(defs) >= {"FLAG_NO_GRAPH", "FLAG_X_SHARD_ONLY", "FLAG_GRAPH_LOOP", "FLAG_DENSE", "FLAG_SPARSE"}
    for name, value in defs.items():
        if name != "FLAG_NONE":
            assert getattr(capi, name) == value, name


def test_pitch():
    assert [capi.pitch_for(d) for d in (1, 4, 5, 123, 128, 784)] == [4, 4, 8, 124, 128, 784]


def test_shard_layout_matches_oracle(oracle):
    for n, p in [(60000, 8), (32561, 8), (10, 4), (7, 1), (1000000, 8), (4000000, 8), (5, 2)]:
        s, d = capi.shard_layout(n, p)
        so, do = oracle.shards(n, p)
        assert s.tolist() == so.tolist() and d.tolist() == do.tolist()
    with pytest.raises(capi.DpsvmError):      # the reference would give the last rank -5 rows here
        capi.shard_layout(9, 8)


def test_merge_candidates_matches_oracle(oracle):
    rng = np.random.default_rng(1)
    for world in (1, 2, 4, 8):
        recv = rng.standard_normal((world, 4)).astype(np.float32)
        recv[:, :2] = rng.integers(0, 1 << 20, (world, 2))
        recv[world // 2, 2] = recv[0, 2]           # a tie: lowest rank wins
        assert capi.merge_candidates(recv) == oracle.merge_candidates(recv)
    empty = np.array([-1, -1, 1e9, -1e9], np.float32)
    assert capi.merge_candidates(empty) == (0, 0, np.float32(1e9), np.float32(-1e9))


def test_alpha_update_matches_oracle(oracle):
    rng = np.random.default_rng(2)
    for _ in range(200):
        eta = float(rng.choice([0.0, 1e-8, 0.3, 1.7, 2.0]))
        yh, yl = int(rng.choice([-1, 1])), int(rng.choice([-1, 1]))
        bh, bl = float(np.float32(rng.normal())), float(np.float32(rng.normal()))
        ah, al = float(np.float32(rng.uniform(0, 10))), float(np.float32(rng.uniform(0, 10)))
        got = capi.alpha_update(eta, yh, yl, bh, bl, ah, al, 10.0)
        want = oracle.alpha_update(eta, yh, yl, bh, bl, ah, al, 10.0)
        assert np.array_equal(np.array(got), np.array(want), equal_nan=True)


def test_packed_keys_order_like_floats():
    rng = np.random.default_rng(3)
    vals = np.concatenate([rng.standard_normal(300).astype(np.float32) * 100,
                           np.array([0.0, -0.0, 1e9, -1e9, 1e-40, -1e-40, 3.0, 3.0], np.float32)])
    idx = rng.integers(0, 1 << 30, vals.shape[0])
    ku = [capi.pack_key_up(v, i) for v, i in zip(vals, idx)]
    kl = [capi.pack_key_low(v, i) for v, i in zip(vals, idx)]
    # min over key_up == argmin f with lowest index among equals; min over key_low == argmax f likewise
    order = np.lexsort((idx, vals))
    assert int(np.argmin(np.array(ku, dtype=np.uint64))) == int(order[0])
    order_hi = np.lexsort((idx, -vals))
    assert int(np.argmin(np.array(kl, dtype=np.uint64))) == int(order_hi[0])
    for v, i, a, b in zip(vals, idx, ku, kl):
        fv, iv = capi.unpack_key_up(a)
        assert fv == v and iv == i
        fv, iv = capi.unpack_key_low(b)
        assert fv == v and iv == i
    # -0.0 and +0.0 compare equal as floats, so they must produce the same key
    assert capi.pack_key_up(-0.0, 7) == capi.pack_key_up(0.0, 7)
    assert capi.pack_key_low(-0.0, 7) == capi.pack_key_low(0.0, 7)
    assert capi.unpack_key_up(capi.KEY_EMPTY) == (np.float32(1e9), 0)
    assert capi.unpack_key_low(capi.KEY_EMPTY) == (np.float32(-1e9), 0)


def test_bad_arguments_are_rejected():
    with pytest.raises(capi.DpsvmError) as e:
        capi.Trainer(100, 0, 1.0, 0.5, 1e-3)
    assert e.value.code == -1
    with pytest.raises(capi.DpsvmError):
        capi.Trainer(100, 8, 1.0, 0.5, 1e-3, shard_rows=60, shard_start=50)


@pytest.mark.skipif(_have_gpu(), reason="checks the no-device behaviour")
def test_no_cpu_fallback_without_a_device():
    with pytest.raises(capi.DpsvmError) as e:
        capi.Trainer(100, 8, 1.0, 0.5, 1e-3)
    assert e.value.code == -2 and "no CUDA device" in str(e.value)


def test_product_does_not_import_the_oracle():
    """Only tests/, bench.py's baseline legs and smoke() may touch oracle/."""
    pkg = os.path.join(ROOT, "dpsvm_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                src = open(os.path.join(dirpath, fn), errors="replace").read()
                assert "oracle" not in src.replace("the oracle", ""), fn + " references oracle/"
