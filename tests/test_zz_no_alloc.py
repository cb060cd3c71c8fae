"""dpsvm_smo_iteration_noalloc -- the entry point for callers that cannot allocate device memory
(a plugin under a framework allocator that owns the device, a thread that must avoid the device-wide
synchronisation of cudaMalloc, a fixed memory budget).  It must do exactly what one iteration of the
trainer context does, with its work area in the library's built-in array instead of cudaMalloc'ed
buffers.

The file name sorts last on purpose: with `pytest -x` every other test has already run by the time
this one can fail.  The test is self-contained, so it also passes alone in a fresh process; and it
makes two calls in a row because the built-in work area keeps the previous call's contents.
"""
import numpy as np
import pytest

from dpsvm_b200 import capi

pytestmark = pytest.mark.gpu


def test_noalloc_iteration_matches_trainer_context(golden):
    import torch
    torch.cuda.set_device(0)
    g = golden("mnist_512")
    x, y = np.ascontiguousarray(g["x"]), np.ascontiguousarray(g["y"])
    n, d = x.shape
    c, gamma, eps = float(g["c"]), float(g["gamma"]), float(g["epsilon"])
    pitch = capi.pitch_for(d)

    # caller-owned state, allocated up front by the caller (here: torch), never by the call under test
    xp = np.zeros((n, pitch), np.float32)
    xp[:, :d] = x
    dx = torch.from_numpy(xp).cuda()
    dy = torch.from_numpy(y).cuda()
    dxsq = torch.empty(n, dtype=torch.float32, device="cuda")
    dalpha = torch.zeros(n, dtype=torch.float32, device="cuda")
    df = torch.empty(n, dtype=torch.float32, device="cuda")
    dkeys = torch.zeros(2, dtype=torch.int64, device="cuda")
    dinfo = torch.zeros(6, dtype=torch.int32, device="cuda")
    capi.row_norms(dx, n, d, pitch, dxsq)
    capi.init_state(dy, n, dalpha, df)
    capi.select(dalpha, dy, df, n, c, 0, dkeys)
    torch.cuda.synchronize()

    with capi.Trainer(n, d, c, gamma, eps, cache_size=0, flags=capi.FLAG_DENSE) as t:      # the existing entry points (dense sweep, as the no-alloc call)
        t.upload(x, y)
        for call in range(2):                                        # two calls in a row
            st = t.train(1)
            a_ref, f_ref, _ = t.download()
            _, keys_ref = t.step1()

            before = torch.cuda.memory_allocated()
            capi.smo_iteration_noalloc(d_x=dx, n=n, d=d, pitch=pitch, d_xsq=dxsq, d_y=dy, d_alpha=dalpha, d_f=df,
                                       d_keys=dkeys, c=c, gamma=gamma, epsilon=eps, index_base=0, d_info=dinfo)
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated() == before

            # same kernels, same inputs: the test suite's tolerance for identical code paths is equality
            assert np.array_equal(dalpha.cpu().numpy(), a_ref), "alpha, call %d" % call
            assert np.array_equal(df.cpu().numpy(), f_ref), "f, call %d" % call
            k = dkeys.cpu().numpy().view(np.uint64)
            assert (int(k[0]), int(k[1])) == keys_ref, "keys, call %d" % call
            info = dinfo.cpu().numpy()
            assert (int(info[0]), int(info[1])) == (st.i_hi, st.i_lo)
            assert info[2:5].view(np.float32).tolist() == [st.b_hi, st.b_lo, st.eta]
            assert int(info[5]) == st.converged


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
