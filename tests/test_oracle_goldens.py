"""The oracle (oracle/smo_oracle.c) against the golden vectors produced by the reference's own
seq.cpp (tests/golden, tools/make_goldens.py).  CPU only.  Bit-exact: same trace, alpha, f, b."""
import numpy as np
import pytest

SMALL = ["mnist_512", "adult_2048", "adult_1000_tight", "odd_d_333x37", "dups_96x16", "tiny_3x5"]


@pytest.mark.parametrize("name", SMALL)
def test_port_reproduces_reference_goldens(name, golden, oracle):
    g = golden(name)
    r = oracle.train(g["x"], g["y"], float(g["c"]), float(g["gamma"]), float(g["epsilon"]), int(g["max_iter"]),
                     kernel_form=0, select_rule=0, nthreads=oracle.max_threads(), trace_cap=int(g["num_iter"]))
    assert r.num_iter == int(g["num_iter"])
    assert np.array_equal(r.trace_idx, g["trace_idx"])
    assert np.array_equal(r.trace_b, g["trace_b"])
    assert np.array_equal(r.alpha, g["alpha"])
    assert np.array_equal(r.f, g["f"])
    assert r.b == g["b"]


@pytest.mark.parametrize("name,k", [("mnist_4096", 60), ("mnist_8192_tight", 40)])
def test_port_prefix_of_large_golden(golden, oracle, name, k):
    """mnist_4096 and mnist_8192_tight take minutes (17 for the latter) on one core: check the first iterations of the
    trace only."""
    g = golden(name)
    r = oracle.train(g["x"], g["y"], float(g["c"]), float(g["gamma"]), float(g["epsilon"]), k,
                     nthreads=oracle.max_threads(), trace_cap=k)
    assert np.array_equal(r.trace_idx, g["trace_idx"][:k])
    assert np.array_equal(r.trace_b, g["trace_b"][:k])


def test_selection_rules_agree_on_goldens(golden, oracle):
    """seq.cpp's set-order scan and the lowest-index rule pick the same pair on these sets
    (the golden files record 0 disagreements); the CUDA path implements the lowest-index rule."""
    for name in SMALL + ["mnist_4096", "mnist_1024_tight", "mnist_8192_tight"]:
        assert int(golden(name)["rule_disagreements"]) == 0


def test_lowest_index_rule_on_ties(oracle):
    alpha = np.zeros(8, np.float32)
    y = np.array([1, 1, -1, -1, 1, -1, 1, -1], np.int32)
    f = np.array([-1, -1, 1, 1, -1, 1, -1, 1], np.float32)
    ih, il, bh, bl = oracle.select(alpha, y, f, 10.0, rule=1)
    assert (ih, il, bh, bl) == (0, 2, -1.0, 1.0)
    ih0, il0, _, _ = oracle.select(alpha, y, f, 10.0, rule=0)
    assert (ih0, il0) == (0, 2)
    # a straddling tie: index 1 is in I0 (0<alpha<c), index 0 in I1; seq scans I0 first, lowest-index takes 0
    alpha[1] = 0.5
    ih1, _, _, _ = oracle.select(alpha, y, f, 10.0, rule=1)
    ih2, _, _, _ = oracle.select(alpha, y, f, 10.0, rule=0)
    assert ih1 == 0 and ih2 == 1


def test_known_answer_gemv(oracle):
    """The only known-answer vector the reference carries: testblas.c:21-48 -> (-1, 3, -3).
    Pushed through the shim's sdot, the primitive every oracle dot product uses."""
    import ctypes as C
    k = dict(np.load(__import__("os").path.join(__import__("conftest").GOLDEN, "kat_gemv3x3.npz")))
    m, x, want = k["m"], k["x"], k["y"]
    # |m_i|^2 via oracle_row_norms and <m_i, x> via the polarisation identity keeps to the exported API
    got = []
    for i in range(3):
        s = oracle.row_norms(np.stack([m[i] + x, m[i] - x]))
        got.append((s[0] - s[1]) / 4)
    assert np.allclose(got, want, atol=1e-5)


def test_empty_sets_use_reference_defaults(oracle):
    # all rows at alpha == c with y == +1: I_up is empty -> index 0, +1e9 (svmTrainMain.cpp:249-252)
    alpha = np.full(4, 2.0, np.float32)
    y = np.ones(4, np.int32)
    f = np.arange(4, dtype=np.float32)
    ih, il, bh, bl = oracle.select(alpha, y, f, 2.0, rule=1)
    assert ih == 0 and bh == np.float32(1e9)
    assert il == 3 and bl == 3.0


def test_shards_and_merge(oracle):
    size, disp = oracle.shards(60000, 8)
    assert size.tolist() == [7500] * 8 and disp.tolist() == [7500 * i for i in range(8)]
    size, disp = oracle.shards(10, 4)   # ceil(10/4)=3 -> 3,3,3,1
    assert size.tolist() == [3, 3, 3, 1]
    recv = np.array([5, 7, -0.5, 0.25, 11, 13, -0.5, 0.75], np.float32)   # tie on f_hi: lowest rank wins
    assert oracle.merge_candidates(recv) == (5, 13, np.float32(-0.5), np.float32(0.75))


def test_lru_replay(oracle):
    hits, hit, line = oracle.lru_replay(np.array([1, 2, 1, 3, 2, 4, 1], np.int32), 2)
    #   1 miss(l0) 2 miss(l1) 1 hit 3 miss evicts 2 (l1) 2 miss evicts 1 (l0) 4 miss evicts 3 (l1) 1 miss evicts 2 (l0)
    assert hit.tolist() == [0, 0, 1, 0, 0, 0, 0] and hits == 1
    assert line.tolist() == [0, 1, 0, 1, 0, 1, 0]


def test_dot_order_restatements(oracle):
    """oracle_dot_dense_order / oracle_dot_sparse_order (the CUDA path's two summation orders): the reference's only
    known-answer vector (testblas.c:21-33, 3x3 gemv -> (-1, 3, -3)) exactly, random rows to fp32 rounding, and
    the sparse order ignores zeros (including -0.0) wherever they sit."""
    k = dict(np.load(__import__("os").path.join(__import__("conftest").GOLDEN, "kat_gemv3x3.npz")))
    for r in range(3):
        assert float(oracle.dot_dense_order(k["m"][r], k["x"])) == float(k["y"][r])
        for lanes in (4, 8, 16, 32):
            assert float(oracle.dot_sparse_order(k["m"][r], k["x"], lanes)) == float(k["y"][r])
    rng = np.random.default_rng(11)
    for d in (1, 3, 5, 37, 123, 128, 520, 784, 1100):
        x = ((rng.random(d) < 0.4) * rng.standard_normal(d)).astype(np.float32)
        h = rng.standard_normal(d).astype(np.float32)
        ref = float(x.astype(np.float64) @ h.astype(np.float64))
        scale = float(np.abs(x.astype(np.float64) * h).sum()) + 1e-30
        assert abs(float(oracle.dot_dense_order(x, h)) - ref) < 4e-7 * scale
        for lanes in (4, 8, 16, 32):
            v = oracle.dot_sparse_order(x, h, lanes)
            assert abs(float(v) - ref) < 4e-7 * scale
            x2 = np.concatenate([np.array([-0.0, 0.0], np.float32), x, np.zeros(7, np.float32)])
            h2 = np.concatenate([np.array([5.0, -3.0], np.float32), h, np.ones(7, np.float32)])
            assert oracle.dot_sparse_order(x2, h2, lanes) == v
