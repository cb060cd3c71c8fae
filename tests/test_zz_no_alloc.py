"""dpsvm_smo_iteration_noalloc -- the entry point for callers that cannot allocate device memory
(a plugin under a framework allocator that owns the device, a thread that must avoid the device-wide
synchronisation of cudaMalloc, a fixed memory budget).  It must do exactly what one iteration of the
trainer context does, with its work area in the library's built-in array instead of cudaMalloc'ed
buffers.

The file name sorts last on purpose: with `pytest -x` every other test has already run by the time
this one can fail.  The tests are self-contained, so they also pass alone in a fresh process; and they
make two (or more) calls in a row because the built-in work area keeps the previous call's contents.
"""
import numpy as np
import pytest

from dpsvm_b200 import capi

pytestmark = pytest.mark.gpu


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


class _CallerState:
    """Everything the call needs, allocated up front by the CALLER (here: torch) -- never by the call under test."""

    def __init__(self, torch, x, y, c, index_base=0, device=0):
        self.torch = torch
        n, d = x.shape
        self.n, self.d, self.c, self.base = n, d, c, index_base
        self.pitch = capi.pitch_for(d)
        xp = np.zeros((n, self.pitch), np.float32)
        xp[:, :d] = x
        dev = "cuda:%d" % device
        self.dx = torch.from_numpy(xp).to(dev)
        self.dy = torch.from_numpy(np.ascontiguousarray(y)).to(dev)
        self.dxsq = torch.empty(n, dtype=torch.float32, device=dev)
        self.dalpha = torch.zeros(n, dtype=torch.float32, device=dev)
        self.df = torch.empty(n, dtype=torch.float32, device=dev)
        self.dkeys = torch.zeros(2, dtype=torch.int64, device=dev)
        self.dinfo = torch.zeros(6, dtype=torch.int32, device=dev)
        capi.row_norms(self.dx, n, d, self.pitch, self.dxsq)
        capi.init_state(self.dy, n, self.dalpha, self.df)
        capi.select(self.dalpha, self.dy, self.df, n, c, index_base, self.dkeys)
        torch.cuda.synchronize(device)

    def iterate(self, gamma, eps):
        capi.smo_iteration_noalloc(d_x=self.dx, n=self.n, d=self.d, pitch=self.pitch, d_xsq=self.dxsq, d_y=self.dy,
                                   d_alpha=self.dalpha, d_f=self.df, d_keys=self.dkeys, c=self.c, gamma=gamma, epsilon=eps,
                                   index_base=self.base, d_info=self.dinfo)

    def keys(self):
        k = self.dkeys.cpu().numpy().view(np.uint64)
        return int(k[0]), int(k[1])


def _compare_with_trainer(torch, x, y, c, gamma, eps, calls, device=0, check_free_memory=False):
    """`calls` iterations through the entry point under test against the same iterations of a trainer context (the
    existing entry points; dense sweep and no cache, as the no-alloc call).  Same kernels, same inputs: the suite's
    tolerance for identical code paths is bit equality."""
    torch.cuda.set_device(device)
    s = _CallerState(torch, x, y, c, device=device)
    n, d = x.shape
    with capi.Trainer(n, d, c, gamma, eps, cache_size=0, device=device, flags=capi.FLAG_DENSE | capi.FLAG_GRAPH_LOOP) as t:
        t.upload(np.ascontiguousarray(x), np.ascontiguousarray(y))
        for call in range(calls):
            st = t.train(1)
            a_ref, f_ref, _ = t.download()
            _, keys_ref = t.step1()

            # "allocates nothing": free device memory as the driver counts it (cudaMemGetInfo -- this sees a cudaMalloc
            # made by ANY library in the process, unlike torch's own allocator statistics) must not move across a call.
            # Checked from the second call on: the very first call loads the kernels' module code.
            torch.cuda.synchronize(device)
            free_before = torch.cuda.mem_get_info(device)[0]
            s.iterate(gamma, eps)
            torch.cuda.synchronize(device)
            if check_free_memory and call > 0:
                assert torch.cuda.mem_get_info(device)[0] == free_before, "free device memory moved across call %d" % call

            assert np.array_equal(s.dalpha.cpu().numpy(), a_ref), "alpha, call %d" % call
            assert np.array_equal(s.df.cpu().numpy(), f_ref), "f, call %d" % call
            assert s.keys() == keys_ref, "keys, call %d" % call
            info = s.dinfo.cpu().numpy()
            assert (int(info[0]), int(info[1])) == (st.i_hi, st.i_lo)
            assert info[2:5].view(np.float32).tolist() == [st.b_hi, st.b_lo, st.eta]
            assert int(info[5]) == st.converged


def test_noalloc_iteration_matches_trainer_context(golden):
    import torch
    g = golden("mnist_512")
    _compare_with_trainer(torch, g["x"], g["y"], float(g["c"]), float(g["gamma"]), float(g["epsilon"]), calls=3,
                          check_free_memory=True)


@pytest.mark.parametrize("d", [1100, 4096])
def test_noalloc_wide_rows(d):
    """Rows wider than 1024 floats take the runtime-length kernels (which set a function attribute on first use);
    4096 floats is the widest row the built-in work area holds."""
    import torch
    rng = np.random.default_rng(d)
    n = 96
    x = (rng.random((n, d), dtype=np.float32) < 0.2).astype(np.float32) * rng.random((n, d), dtype=np.float32)
    y = np.where(rng.random(n) < 0.5, 1, -1).astype(np.int32)
    _compare_with_trainer(torch, x, y, 4.0, 0.02, 1e-3, calls=3, check_free_memory=True)


@pytest.mark.skipif(_ngpu() < 2, reason="needs a second GPU")
def test_noalloc_on_a_second_device(golden):
    """The work area exists once per GPU and its address is looked up per device: calls on device 1, then device 0
    again, each against a trainer on the same device."""
    import torch
    g = golden("odd_d_333x37")
    for dev in (1, 0, 1):
        _compare_with_trainer(torch, g["x"], g["y"], float(g["c"]), float(g["gamma"]), float(g["epsilon"]), calls=2, device=dev)
    torch.cuda.set_device(0)


def test_noalloc_one_class_rows_with_a_nonzero_index_base():
    """Rows of a single class leave one I-set empty.  An empty set stands for the reference's default "row 0"
    (svmTrainMain.cpp:249-252); with index_base > 0 global row 0 is not among the caller's rows, and the call must then
    use the first row it was given instead of reaching outside the caller's arrays."""
    import torch
    torch.cuda.set_device(0)
    rng = np.random.default_rng(3)
    n, d = 200, 24
    x = rng.random((n, d), dtype=np.float32)
    y = np.ones(n, np.int32)
    res = []
    for base in (0, 100000):
        s = _CallerState(torch, x, y, 2.0, index_base=base)
        k0 = s.keys()
        assert k0[1] == capi.KEY_EMPTY and capi.unpack_key_up(k0[0]) == (np.float32(-1.0), base)
        s.iterate(0.3, 1e-3)
        torch.cuda.synchronize()
        info = s.dinfo.cpu().numpy()
        res.append((s.dalpha.cpu().numpy(), s.df.cpu().numpy(), int(info[0]) - base, int(info[1]) - base,
                    info[2:5].view(np.float32).tolist(), int(info[5])))
        assert np.isfinite(res[-1][1]).all()
    assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1]) and res[0][2:] == res[1][2:]
    assert res[0][2:4] == (0, 0) and res[0][5] == 1          # pair (first row, first row); the empty set ends the loop


def test_noalloc_rejects_rows_wider_than_the_work_area():
    import torch
    torch.cuda.set_device(0)
    pitch = capi.NOALLOC_MAX_PITCH + 4
    z = torch.zeros(2 * pitch, dtype=torch.float32, device="cuda")
    k = torch.zeros(2, dtype=torch.int64, device="cuda")
    with pytest.raises(capi.DpsvmError) as e:
        capi.smo_iteration_noalloc(d_x=z, n=2, d=pitch, pitch=pitch, d_xsq=z, d_y=k, d_alpha=z, d_f=z, d_keys=k,
                                   c=1.0, gamma=0.1, epsilon=1e-3, index_base=0, d_info=None)
    assert e.value.code == -6 and "DPSVM_NOALLOC_MAX_PITCH" in str(e.value)


def test_noalloc_rejects_a_misaligned_row_pointer():
    import torch
    torch.cuda.set_device(0)
    z = torch.zeros(64, dtype=torch.float32, device="cuda")
    k = torch.zeros(2, dtype=torch.int64, device="cuda")
    with pytest.raises(capi.DpsvmError) as e:
        capi.smo_iteration_noalloc(d_x=z.data_ptr() + 4, n=2, d=8, pitch=8, d_xsq=z, d_y=k, d_alpha=z, d_f=z, d_keys=k,
                                   c=1.0, gamma=0.1, epsilon=1e-3, index_base=0, d_info=None)
    assert e.value.code == -1 and "16-byte" in str(e.value)
