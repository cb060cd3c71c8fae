"""ctypes binding of libdpsvm.so (include/dpsvm_cuda.h) -- glue for tests and bench.py.

This is a thin binding, not an implementation: every compute call goes to the CUDA library, and
loading fails loudly when the library has not been built (``python -c 'import __graft_entry__ as g;
g.build()'`` or ``make lib``).  There is no CPU fallback anywhere in this package.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdpsvm.so")

OK = 0
FLAG_NO_GRAPH = 1
FLAG_X_SHARD_ONLY = 2
FLAG_GRAPH_LOOP = 4
FLAG_DENSE = 8
FLAG_SPARSE = 16
FLAG_X_MIRROR = 32
KEY_EMPTY = 0xFFFFFFFFFFFFFFFF
COMM_ID_BYTES = 128


class DpsvmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("dpsvm error %d: %s" % (code, msg))
        self.code = code


class Params(C.Structure):
    _fields_ = [("num_attributes", C.c_int), ("num_train_data", C.c_int), ("c", C.c_float), ("gamma", C.c_float),
                ("epsilon", C.c_float), ("max_iter", C.c_int), ("cache_size", C.c_int), ("shard_rows", C.c_int),
                ("shard_start", C.c_int), ("device", C.c_int), ("flags", C.c_int)]


class TrainStatus(C.Structure):
    _fields_ = [("num_iter", C.c_int), ("converged", C.c_int), ("b_hi", C.c_float), ("b_lo", C.c_float),
                ("b", C.c_float), ("cache_hits", C.c_longlong), ("cache_lookups", C.c_longlong),
                ("last_chunk_ms", C.c_float), ("i_hi", C.c_int), ("i_lo", C.c_int), ("eta", C.c_float)]


class PassArgs(C.Structure):
    _fields_ = [("d_x", C.c_void_p), ("n", C.c_int), ("d", C.c_int), ("pitch", C.c_int),
                ("d_xsq", C.c_void_p), ("d_y", C.c_void_p), ("d_alpha", C.c_void_p), ("d_f", C.c_void_p),
                ("d_x_hi", C.c_void_p), ("d_x_lo", C.c_void_p),
                ("xsq_hi", C.c_float), ("xsq_lo", C.c_float), ("delta_hi", C.c_float), ("delta_lo", C.c_float),
                ("gamma", C.c_float), ("c", C.c_float), ("index_base", C.c_int),
                ("d_keys", C.c_void_p), ("d_line_hi", C.c_void_p), ("d_line_lo", C.c_void_p), ("variant", C.c_int)]


class IterationInfo(C.Structure):
    _fields_ = [("i_hi", C.c_int32), ("i_lo", C.c_int32), ("b_hi", C.c_float), ("b_lo", C.c_float),
                ("eta", C.c_float), ("converged", C.c_int32)]


class IterationArgs(C.Structure):
    _fields_ = [("d_x", C.c_void_p), ("n", C.c_int), ("d", C.c_int), ("pitch", C.c_int),
                ("d_xsq", C.c_void_p), ("d_y", C.c_void_p), ("d_alpha", C.c_void_p), ("d_f", C.c_void_p),
                ("d_keys", C.c_void_p), ("c", C.c_float), ("gamma", C.c_float), ("epsilon", C.c_float),
                ("index_base", C.c_int), ("d_info", C.c_void_p)]


NOALLOC_MAX_PITCH = 4096

# every symbol include/dpsvm_cuda.h declares (tests check the library exports all of them)
EXPORTS = [
    "dpsvm_last_error", "dpsvm_abi_version", "dpsvm_shard_layout", "dpsvm_merge_candidates", "dpsvm_alpha_update",
    "dpsvm_pitch_for", "dpsvm_pass_variant", "dpsvm_pack_key_up", "dpsvm_pack_key_low", "dpsvm_unpack_key_up", "dpsvm_unpack_key_low",
    "dpsvm_row_norms", "dpsvm_init_state", "dpsvm_select", "dpsvm_smo_pass", "dpsvm_create", "dpsvm_destroy",
    "dpsvm_upload", "dpsvm_upload_device", "dpsvm_reset", "dpsvm_step1", "dpsvm_step2", "dpsvm_train",
    "dpsvm_train_async", "dpsvm_loop_info", "dpsvm_layout_info", "dpsvm_sync", "dpsvm_stream", "dpsvm_download", "dpsvm_device_state", "dpsvm_cache_lookup",
    "dpsvm_predict", "dpsvm_comm_unique_id", "dpsvm_comm_init", "dpsvm_comm_transport", "dpsvm_smo_iteration_noalloc",
]

_lib = None


def lib():
    """Load libdpsvm.so (once).  Raises if it is missing: the product has no other path."""
    global _lib
    if _lib is None:
        path = os.environ.get("DPSVM_LIB", LIB_PATH)          # A/B timing of two builds; normally the in-tree library
        if not os.path.exists(path):
            raise FileNotFoundError(path + " not built -- run __graft_entry__.build() or `make lib`")
        L = C.CDLL(path)
        L.dpsvm_last_error.restype = C.c_char_p
        L.dpsvm_pack_key_up.restype = C.c_uint64
        L.dpsvm_pack_key_up.argtypes = [C.c_float, C.c_int32]
        L.dpsvm_pack_key_low.restype = C.c_uint64
        L.dpsvm_pack_key_low.argtypes = [C.c_float, C.c_int32]
        L.dpsvm_unpack_key_up.argtypes = [C.c_uint64, C.POINTER(C.c_float), C.POINTER(C.c_int32)]
        L.dpsvm_unpack_key_low.argtypes = [C.c_uint64, C.POINTER(C.c_float), C.POINTER(C.c_int32)]
        L.dpsvm_alpha_update.argtypes = [C.c_float, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                                         C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.dpsvm_row_norms.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.dpsvm_init_state.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.dpsvm_select.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
        L.dpsvm_smo_pass.argtypes = [C.POINTER(PassArgs), C.c_void_p]
        L.dpsvm_smo_iteration_noalloc.argtypes = [C.POINTER(IterationArgs), C.c_void_p]
        L.dpsvm_create.argtypes = [C.POINTER(Params), C.POINTER(C.c_void_p)]
        L.dpsvm_destroy.argtypes = [C.c_void_p]
        L.dpsvm_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.dpsvm_upload_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.dpsvm_reset.argtypes = [C.c_void_p]
        L.dpsvm_layout_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
        L.dpsvm_step1.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_uint64)]
        L.dpsvm_step2.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.dpsvm_train.argtypes = [C.c_void_p, C.c_int, C.POINTER(TrainStatus)]
        L.dpsvm_train_async.argtypes = [C.c_void_p, C.c_int]
        L.dpsvm_sync.argtypes = [C.c_void_p, C.POINTER(TrainStatus)]
        L.dpsvm_loop_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.dpsvm_stream.restype = C.c_void_p
        L.dpsvm_stream.argtypes = [C.c_void_p]
        L.dpsvm_download.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.dpsvm_device_state.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        L.dpsvm_cache_lookup.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int)]
        L.dpsvm_predict.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p, C.c_int,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_float)]
        L.dpsvm_comm_unique_id.argtypes = [C.c_void_p]
        L.dpsvm_comm_init.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.dpsvm_comm_transport.argtypes = [C.c_void_p]
        L.dpsvm_shard_layout.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.dpsvm_merge_candidates.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                             C.POINTER(C.c_float), C.POINTER(C.c_float)]
        _lib = L
    return _lib


def check(rc):
    if rc != OK:
        raise DpsvmError(rc, lib().dpsvm_last_error().decode(errors="replace"))


def _ptr(a):
    """Host numpy array, torch tensor (host or device), raw int address or None -> void*."""
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    raise TypeError(type(a))


# ---- host-only helpers -----------------------------------------------------------------------
def pitch_for(d: int) -> int:
    return lib().dpsvm_pitch_for(d)


def pass_variant(n: int, pitch: int, parity: int) -> int:
    return lib().dpsvm_pass_variant(n, pitch, parity)


def shard_layout(n_total: int, world: int):
    size = np.zeros(world, np.int32)
    disp = np.zeros(world, np.int32)
    check(lib().dpsvm_shard_layout(n_total, world, _ptr(size), _ptr(disp)))
    return size, disp


def merge_candidates(recv):
    recv = np.ascontiguousarray(recv, np.float32).reshape(-1)
    ih, il, bh, bl = C.c_int(), C.c_int(), C.c_float(), C.c_float()
    check(lib().dpsvm_merge_candidates(_ptr(recv), recv.shape[0] // 4, C.byref(ih), C.byref(il), C.byref(bh), C.byref(bl)))
    return ih.value, il.value, np.float32(bh.value), np.float32(bl.value)


def alpha_update(eta, y_hi, y_lo, b_hi, b_lo, a_hi, a_lo, c):
    ah, al = C.c_float(), C.c_float()
    check(lib().dpsvm_alpha_update(eta, y_hi, y_lo, b_hi, b_lo, a_hi, a_lo, c, C.byref(ah), C.byref(al)))
    return np.float32(ah.value), np.float32(al.value)


def pack_key_up(f, idx):
    return int(lib().dpsvm_pack_key_up(float(f), int(idx)))


def pack_key_low(f, idx):
    return int(lib().dpsvm_pack_key_low(float(f), int(idx)))


def unpack_key_up(k):
    f, i = C.c_float(), C.c_int32()
    lib().dpsvm_unpack_key_up(C.c_uint64(k), C.byref(f), C.byref(i))
    return np.float32(f.value), i.value


def unpack_key_low(k):
    f, i = C.c_float(), C.c_int32()
    lib().dpsvm_unpack_key_low(C.c_uint64(k), C.byref(f), C.byref(i))
    return np.float32(f.value), i.value


# ---- stateless launchers (device pointers: torch CUDA tensors or raw addresses) -----------------
def row_norms(d_x, n, d, pitch, d_xsq, stream=None):
    check(lib().dpsvm_row_norms(_ptr(d_x), n, d, pitch, _ptr(d_xsq), _ptr(stream)))


def init_state(d_y, n, d_alpha, d_f, stream=None):
    check(lib().dpsvm_init_state(_ptr(d_y), n, _ptr(d_alpha), _ptr(d_f), _ptr(stream)))


def select(d_alpha, d_y, d_f, n, c, index_base, d_keys, stream=None):
    check(lib().dpsvm_select(_ptr(d_alpha), _ptr(d_y), _ptr(d_f), n, c, index_base, _ptr(d_keys), _ptr(stream)))


def smo_pass(stream=None, **kw):
    a = PassArgs()
    for k, v in kw.items():
        if k.startswith("d_"):
            p = _ptr(v)
            setattr(a, k, p.value if p is not None else None)
        else:
            setattr(a, k, v)
    check(lib().dpsvm_smo_pass(C.byref(a), _ptr(stream)))


def smo_iteration_noalloc(stream=None, **kw):
    """One SMO iteration on caller-owned device arrays with no allocation (dpsvm_smo_iteration_noalloc)."""
    a = IterationArgs()
    for k, v in kw.items():
        if k.startswith("d_"):
            p = _ptr(v)
            setattr(a, k, p.value if p is not None else None)
        else:
            setattr(a, k, v)
    check(lib().dpsvm_smo_iteration_noalloc(C.byref(a), _ptr(stream)))


# ---- trainer context --------------------------------------------------------------------------
class Trainer:
    """One rank's trainer (dpsvm_ctx).  Mirrors SvmTrain's life cycle: create -> upload (setup) ->
    step1/step2 or train -> download / predict -> close."""

    def __init__(self, n_total, d, c, gamma, epsilon, max_iter=150000, cache_size=10, shard_rows=None,
                 shard_start=0, device=-1, flags=0):
        self.params = Params(d, n_total, c, gamma, epsilon, max_iter, cache_size,
                             n_total if shard_rows is None else shard_rows, shard_start, device, flags)
        self._h = C.c_void_p()
        check(lib().dpsvm_create(C.byref(self.params), C.byref(self._h)))
        self.n_total, self.d = n_total, d
        self.shard_rows = self.params.shard_rows

    def close(self):
        if self._h:
            lib().dpsvm_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, x, y):
        """x: host float32 [rows, d] (numpy or pinned torch tensor), y: int32 [rows]."""
        self._keep = (x, y)
        check(lib().dpsvm_upload(self._h, _ptr(x), _ptr(y)))
        self._keep = None

    def upload_device(self, d_x, pitch, d_y):
        check(lib().dpsvm_upload_device(self._h, _ptr(d_x), pitch, _ptr(d_y)))

    def reset(self):
        check(lib().dpsvm_reset(self._h))

    def step1(self):
        rv = (C.c_float * 4)()
        keys = (C.c_uint64 * 2)()
        check(lib().dpsvm_step1(self._h, rv, keys))
        return np.array(rv[:], np.float32), (int(keys[0]), int(keys[1]))

    def step2(self, i_hi, i_lo, a_hi, a_lo, x_hi=None, x_lo=None):
        check(lib().dpsvm_step2(self._h, int(i_hi), int(i_lo), float(a_hi), float(a_lo), _ptr(x_hi), _ptr(x_lo)))

    def train(self, max_new_iter=0) -> TrainStatus:
        st = TrainStatus()
        check(lib().dpsvm_train(self._h, max_new_iter, C.byref(st)))
        return st

    def train_async(self, iters):
        check(lib().dpsvm_train_async(self._h, iters))

    def sync(self) -> TrainStatus:
        st = TrainStatus()
        check(lib().dpsvm_sync(self._h, C.byref(st)))
        return st

    def loop_info(self):
        """(resident: bool, chunk: int) -- see dpsvm_loop_info."""
        r, k = C.c_int(), C.c_int()
        check(lib().dpsvm_loop_info(self._h, C.byref(r), C.byref(k)))
        return bool(r.value), k.value

    def layout_info(self):
        """(lanes, nnz, sweep_bytes) -- lanes 0 = dense rows, else the sparse stream; see dpsvm_layout_info."""
        g, nnz, b = C.c_int(), C.c_longlong(), C.c_longlong()
        check(lib().dpsvm_layout_info(self._h, C.byref(g), C.byref(nnz), C.byref(b)))
        return g.value, nnz.value, b.value

    @property
    def stream(self):
        return lib().dpsvm_stream(self._h)

    def download(self, alpha=True, f=True, xsq=False):
        n = self.shard_rows
        a = np.empty(n, np.float32) if alpha else None
        ff = np.empty(n, np.float32) if f else None
        xs = np.empty(n, np.float32) if xsq else None
        check(lib().dpsvm_download(self._h, _ptr(a), _ptr(ff), _ptr(xs)))
        return a, ff, xs

    def device_state(self):
        x, xsq, y, alpha, f = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        pitch = C.c_int()
        check(lib().dpsvm_device_state(self._h, C.byref(x), C.byref(pitch), C.byref(xsq), C.byref(y),
                                       C.byref(alpha), C.byref(f)))
        return dict(x=x.value, pitch=pitch.value, xsq=xsq.value, y=y.value, alpha=alpha.value, f=f.value)

    def cache_lookup(self, key):
        line, hit = C.c_void_p(), C.c_int()
        check(lib().dpsvm_cache_lookup(self._h, key, C.byref(line), C.byref(hit)))
        return line.value, bool(hit.value)

    def predict(self, alpha_full, y_full, b, subtract_b, t=None, labels=None):
        alpha_full = np.ascontiguousarray(alpha_full, np.float32)
        y_full = np.ascontiguousarray(y_full, np.int32)
        if t is not None:
            t = np.ascontiguousarray(t, np.float32)
            n_test = t.shape[0]
        else:
            n_test = self.n_total
        if labels is not None:
            labels = np.ascontiguousarray(labels, np.int32)
        dual = np.empty(n_test, np.float32)
        pred = np.empty(n_test, np.int32)
        nsv, acc = C.c_int(), C.c_float()
        check(lib().dpsvm_predict(self._h, _ptr(alpha_full), _ptr(y_full), b, int(subtract_b), _ptr(t), n_test,
                                  _ptr(labels), _ptr(dual), _ptr(pred), C.byref(nsv), C.byref(acc)))
        return dual, pred, nsv.value, acc.value

    def comm_init(self, id_bytes: bytes, rank: int, world: int):
        buf = (C.c_ubyte * COMM_ID_BYTES).from_buffer_copy(id_bytes)
        check(lib().dpsvm_comm_init(self._h, buf, rank, world))

    @property
    def comm_transport(self) -> str:
        return {0: "none", 1: "nccl", 2: "p2p", 3: "p2p-ll"}[lib().dpsvm_comm_transport(self._h)]


def comm_unique_id() -> bytes:
    buf = (C.c_ubyte * COMM_ID_BYTES)()
    check(lib().dpsvm_comm_unique_id(buf))
    return bytes(buf)
