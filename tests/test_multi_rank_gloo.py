"""Host-side logic of the N > 1 path on CPU: two processes over gloo run the reference's sharded loop
structure (svmTrainMain.cpp:235-310) -- contiguous row shards (dpsvm_shard_layout), per-rank local
selection, an all-gather of the 4-float candidates, the global merge (dpsvm_merge_candidates), the
redundant alpha update (dpsvm_alpha_update) -- with the oracle standing in for the per-shard device
work, and must reproduce the single-process oracle trajectory exactly."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, iters, out_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from dpsvm_b200 import capi, datagen
    from oracle import pyoracle as po
    dist.init_process_group("gloo", init_method="file://" + os.path.join(out_dir, "rdzv"), rank=rank, world_size=world)
    n, d, c, gamma = 301, 24, 10.0, 0.25
    x, y = datagen.mnist_like(n, d, seed=33)
    size, disp = capi.shard_layout(n, world)
    lo, hi = int(disp[rank]), int(disp[rank] + size[rank])
    xs, ys = x[lo:hi], y[lo:hi]
    alpha = np.zeros(n, np.float32)                 # replicated, as the reference keeps it (svmTrainMain.cpp:223)
    f = -ys.astype(np.float32)                      # local shard only
    xsq = po.row_norms(x)
    trace = []
    for _ in range(iters):
        ih, il, bh, bl = po.select(alpha[lo:hi], ys, f, c, rule=1, idx_base=lo)
        rv = torch.tensor([ih, il, bh, bl], dtype=torch.float32)       # indices as floats, svmTrain.cu:478-481
        recv = [torch.zeros(4) for _ in range(world)]
        dist.all_gather(recv, rv)                                     # svmTrainMain.cpp:244
        i_hi, i_lo, b_hi, b_lo = capi.merge_candidates(torch.stack(recv).numpy())
        eta = po.rbf(x[i_hi], x[i_hi], gamma, 1) + po.rbf(x[i_lo], x[i_lo], gamma, 1) - 2 * po.rbf(x[i_lo], x[i_hi], gamma, 1)
        a_hi, a_lo = capi.alpha_update(float(eta), int(y[i_hi]), int(y[i_lo]), float(b_hi), float(b_lo),
                                       float(alpha[i_hi]), float(alpha[i_lo]), c)
        d_hi = np.float32((a_hi - alpha[i_hi]) * y[i_hi])
        d_lo = np.float32((a_lo - alpha[i_lo]) * y[i_lo])
        alpha[i_lo] = a_lo
        alpha[i_hi] = a_hi
        # local f update in the GPU path's kernel form (svmTrain.cu:98-137)
        dot_hi = np.array([np.float32(np.dot(r.astype(np.float64), x[i_hi].astype(np.float64))) for r in xs], np.float32)
        dot_lo = np.array([np.float32(np.dot(r.astype(np.float64), x[i_lo].astype(np.float64))) for r in xs], np.float32)
        k_hi = np.exp(np.float32(-gamma) * (xsq[lo:hi] + xsq[i_hi] - 2 * dot_hi)).astype(np.float32)
        k_lo = np.exp(np.float32(-gamma) * (xsq[lo:hi] + xsq[i_lo] - 2 * dot_lo)).astype(np.float32)
        f = (f + (d_hi * k_hi + d_lo * k_lo)).astype(np.float32)
        trace.append((i_hi, i_lo))
    np.save(os.path.join(out_dir, "trace%d.npy" % rank), np.array(trace, np.int32))
    np.save(os.path.join(out_dir, "alpha%d.npy" % rank), alpha)
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_sharded_loop_matches_single_process(tmp_path, oracle):
    from dpsvm_b200 import datagen
    world, iters = 2, 40
    mp.spawn(_worker, args=(world, _free_port(), iters, str(tmp_path)), nprocs=world, join=True)
    t0, t1 = np.load(tmp_path / "trace0.npy"), np.load(tmp_path / "trace1.npy")
    a0, a1 = np.load(tmp_path / "alpha0.npy"), np.load(tmp_path / "alpha1.npy")
    assert np.array_equal(t0, t1) and np.array_equal(a0, a1)          # every rank takes the same decisions
    x, y = datagen.mnist_like(301, 24, seed=33)
    ref = oracle.train(x, y, 10.0, 0.25, 1e-3, iters, kernel_form=1, select_rule=1, trace_cap=iters)
    assert np.array_equal(t0, ref.trace_idx)                           # and they are the single-process ones
    assert np.abs(a0 - ref.alpha).max() < 1e-5
