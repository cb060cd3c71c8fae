"""ctypes bindings for the CHECKER libraries (test infrastructure, not product code).

* ``libsmo_oracle.so``   -- oracle/smo_oracle.c, the plain-C restatement ("port").
* ``_ref/libseqref.so``  -- the reference's own seq.cpp compiled from /root/reference through
                            oracle/ref_harness.cpp ("reference"); present when it was built in a
                            container that has /root/reference (the .so travels to the GPU box).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_PORT = os.path.join(HERE, "libsmo_oracle.so")
_REF = os.path.join(HERE, "_ref", "libseqref.so")

_fp = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def build(verbose: bool = False) -> None:
    """(Re)build the checker libraries with oracle/Makefile (builds _ref only if /root/reference exists)."""
    r = subprocess.run(["make", "-C", HERE], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout)


def _opt(a, dtype):
    if a is None:
        return None
    assert a.dtype == dtype and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


_port = None
_ref = None


def port():
    global _port
    if _port is None:
        if not os.path.exists(_PORT):
            build()
        _port = C.CDLL(_PORT)
        _port.oracle_rbf.restype = C.c_float
        _port.oracle_rbf.argtypes = [_fp, _fp, C.c_int, C.c_float, C.c_int]
    return _port


def have_ref() -> bool:
    return os.path.exists(_REF)


def ref():
    global _ref
    if _ref is None:
        if not os.path.exists(_REF):
            raise FileNotFoundError(_REF + " (built only where /root/reference exists)")
        _ref = C.CDLL(_REF)
        _ref.seqref_rbf.restype = C.c_float
        _ref.seqref_rbf.argtypes = [_fp, _fp, C.c_int, C.c_float]
    return _ref


class TrainResult(dict):
    __getattr__ = dict.__getitem__


def _prep(x, y):
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.ascontiguousarray(y, dtype=np.int32)
    assert x.ndim == 2 and y.shape == (x.shape[0],)
    return x, y


def train(x, y, c, gamma, epsilon, max_iter=150000, kernel_form=0, select_rule=0,
          alpha0=None, f0=None, nthreads=1, trace_cap=0) -> TrainResult:
    """oracle_smo_train (the port).  Returns alpha, f, num_iter, b_hi, b_lo, b, trace_idx, trace_b,
    rule_disagreements."""
    x, y = _prep(x, y)
    n, d = x.shape
    warm = alpha0 is not None
    alpha = np.array(alpha0, dtype=np.float32) if warm else np.zeros(n, np.float32)
    f = np.array(f0, dtype=np.float32) if warm else np.zeros(n, np.float32)
    tidx = np.zeros((max(trace_cap, 1), 2), np.int32)
    tb = np.zeros((max(trace_cap, 1), 2), np.float32)
    it, bh, bl, dis = C.c_int(0), C.c_float(0), C.c_float(0), C.c_int(0)
    port().oracle_smo_train(
        _opt(x, np.float32), _opt(y, np.int32), C.c_int(n), C.c_int(d),
        C.c_float(c), C.c_float(gamma), C.c_float(epsilon), C.c_int(max_iter),
        C.c_int(kernel_form), C.c_int(select_rule), C.c_int(1 if warm else 0), C.c_int(nthreads),
        _opt(alpha, np.float32), _opt(f, np.float32),
        _opt(tidx, np.int32), _opt(tb, np.float32), C.c_int(trace_cap),
        C.byref(it), C.byref(bh), C.byref(bl), C.byref(dis))
    k = min(it.value, trace_cap)
    return TrainResult(alpha=alpha, f=f, num_iter=it.value, b_hi=np.float32(bh.value),
                       b_lo=np.float32(bl.value),
                       b=np.float32((np.float32(bl.value) + np.float32(bh.value)) / np.float32(2)),
                       trace_idx=tidx[:k], trace_b=tb[:k], rule_disagreements=dis.value)


def ref_train(x, y, c, gamma, epsilon, max_iter=150000, trace_cap=0) -> TrainResult:
    """seqref_train: the reference's own functions (seq.cpp) driven by ref_harness.cpp."""
    x, y = _prep(x, y)
    n, d = x.shape
    alpha = np.zeros(n, np.float32)
    f = np.zeros(n, np.float32)
    tidx = np.zeros((max(trace_cap, 1), 2), np.int32)
    tb = np.zeros((max(trace_cap, 1), 2), np.float32)
    it, bh, bl, dis = C.c_int(0), C.c_float(0), C.c_float(0), C.c_int(0)
    ref().seqref_train(
        _opt(x, np.float32), _opt(y, np.int32), C.c_int(n), C.c_int(d),
        C.c_float(c), C.c_float(gamma), C.c_float(epsilon), C.c_int(max_iter),
        _opt(alpha, np.float32), _opt(f, np.float32),
        _opt(tidx, np.int32), _opt(tb, np.float32), C.c_int(trace_cap),
        C.byref(it), C.byref(bh), C.byref(bl), C.byref(dis))
    k = min(it.value, trace_cap)
    return TrainResult(alpha=alpha, f=f, num_iter=it.value, b_hi=np.float32(bh.value),
                       b_lo=np.float32(bl.value),
                       b=np.float32((np.float32(bl.value) + np.float32(bh.value)) / np.float32(2)),
                       trace_idx=tidx[:k], trace_b=tb[:k], rule_disagreements=dis.value)


def select(alpha, y, f, c, rule=1, idx_base=0):
    """oracle_select -> (i_hi, i_lo, b_hi, b_lo)."""
    alpha = np.ascontiguousarray(alpha, np.float32)
    y = np.ascontiguousarray(y, np.int32)
    f = np.ascontiguousarray(f, np.float32)
    ih, il, bh, bl = C.c_int(0), C.c_int(0), C.c_float(0), C.c_float(0)
    port().oracle_select(_opt(alpha, np.float32), _opt(y, np.int32), _opt(f, np.float32),
                         C.c_int(alpha.shape[0]), C.c_float(c), C.c_int(rule), C.c_int(idx_base),
                         C.byref(ih), C.byref(il), C.byref(bh), C.byref(bl))
    return ih.value, il.value, np.float32(bh.value), np.float32(bl.value)


def rbf(x1, x2, gamma, kernel_form=0) -> np.float32:
    x1 = np.ascontiguousarray(x1, np.float32)
    x2 = np.ascontiguousarray(x2, np.float32)
    return np.float32(port().oracle_rbf(x1, x2, x1.shape[0], C.c_float(gamma), kernel_form))


def ref_rbf(x1, x2, gamma) -> np.float32:
    x1 = np.ascontiguousarray(x1, np.float32)
    x2 = np.ascontiguousarray(x2, np.float32)
    return np.float32(ref().seqref_rbf(x1, x2, x1.shape[0], C.c_float(gamma)))


def row_norms(x):
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros(x.shape[0], np.float32)
    port().oracle_row_norms(_opt(x, np.float32), C.c_int(x.shape[0]), C.c_int(x.shape[1]), _opt(out, np.float32))
    return out


def alpha_update(eta, y_hi, y_lo, b_hi, b_lo, a_hi, a_lo, c):
    ah, al = C.c_float(0), C.c_float(0)
    port().oracle_alpha_update(C.c_float(eta), C.c_int(y_hi), C.c_int(y_lo), C.c_float(b_hi), C.c_float(b_lo),
                               C.c_float(a_hi), C.c_float(a_lo), C.c_float(c), C.byref(ah), C.byref(al))
    return np.float32(ah.value), np.float32(al.value)


def predict(x, y, alpha, gamma, b, t, convention=0):
    """oracle_predict -> (dual float32[n_test], labels int32[n_test])."""
    x, y = _prep(x, y)
    alpha = np.ascontiguousarray(alpha, np.float32)
    t = np.ascontiguousarray(t, np.float32)
    dual = np.zeros(t.shape[0], np.float32)
    lab = np.zeros(t.shape[0], np.int32)
    port().oracle_predict(_opt(x, np.float32), _opt(y, np.int32), _opt(alpha, np.float32),
                          C.c_int(x.shape[0]), C.c_int(x.shape[1]), C.c_float(gamma), C.c_float(b),
                          _opt(t, np.float32), C.c_int(t.shape[0]), C.c_int(convention),
                          _opt(dual, np.float32), _opt(lab, np.int32))
    return dual, lab


def ref_predict(x, y, alpha, gamma, t):
    x, y = _prep(x, y)
    alpha = np.ascontiguousarray(alpha, np.float32)
    t = np.ascontiguousarray(t, np.float32)
    dual = np.zeros(t.shape[0], np.float32)
    lab = np.zeros(t.shape[0], np.int32)
    ref().seqref_predict(_opt(x, np.float32), _opt(y, np.int32), _opt(alpha, np.float32),
                         C.c_int(x.shape[0]), C.c_int(x.shape[1]), C.c_float(gamma),
                         _opt(t, np.float32), C.c_int(t.shape[0]), _opt(dual, np.float32), _opt(lab, np.int32))
    return dual, lab


def shards(n_total, world):
    size = np.zeros(world, np.int32)
    disp = np.zeros(world, np.int32)
    port().oracle_shards(C.c_int(n_total), C.c_int(world), _opt(size, np.int32), _opt(disp, np.int32))
    return size, disp


def merge_candidates(recv):
    recv = np.ascontiguousarray(recv, np.float32).reshape(-1)
    ih, il, bh, bl = C.c_int(0), C.c_int(0), C.c_float(0), C.c_float(0)
    port().oracle_merge_candidates(_opt(recv, np.float32), C.c_int(recv.shape[0] // 4),
                                   C.byref(ih), C.byref(il), C.byref(bh), C.byref(bl))
    return ih.value, il.value, np.float32(bh.value), np.float32(bl.value)


def lru_replay(keys, capacity):
    keys = np.ascontiguousarray(keys, np.int32)
    hit = np.zeros(keys.shape[0], np.uint8)
    line = np.zeros(keys.shape[0], np.int32)
    hits = port().oracle_lru_replay(_opt(keys, np.int32), C.c_int(keys.shape[0]), C.c_int(capacity),
                                    hit.ctypes.data_as(C.c_void_p), _opt(line, np.int32))
    return hits, hit, line


def dot_dense_order(x_row, h) -> np.float32:
    """<x_row, h> summed in the dense kernels' order (oracle_dot_dense_order)."""
    x_row, h = np.ascontiguousarray(x_row, np.float32), np.ascontiguousarray(h, np.float32)
    fn = port().oracle_dot_dense_order
    fn.restype = C.c_float
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    return np.float32(fn(x_row.ctypes.data, h.ctypes.data, x_row.shape[0]))


def dot_sparse_order(x_row, h, lanes) -> np.float32:
    """<x_row, h> summed in the sparse sweep's order with `lanes` lanes per row (oracle_dot_sparse_order)."""
    x_row, h = np.ascontiguousarray(x_row, np.float32), np.ascontiguousarray(h, np.float32)
    fn = port().oracle_dot_sparse_order
    fn.restype = C.c_float
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    return np.float32(fn(x_row.ctypes.data, h.ctypes.data, x_row.shape[0], lanes))


def max_threads() -> int:
    return int(port().oracle_max_threads())
