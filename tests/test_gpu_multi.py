"""Multi-GPU parity (needs >= 2 GPUs: `gpurun --gpus 2`): the sharded, in-graph candidate exchange
must reproduce the single-GPU run bit for bit -- per-row arithmetic does not depend on the sharding,
and every rank takes the same decisions from the same gathered packages."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, cache, exchange, out_dir, layout="auto"):
    graph_loop = exchange.endswith("+graph")
    exchange = exchange.split("+")[0]
    mirror = exchange == "p2p-ll"          # X mirrored on every rank, candidates as flagged words
    if mirror:
        exchange = "p2p"
    os.environ["DPSVM_LAYOUT"] = layout
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    sys.path.insert(0, os.path.join(root, "tests"))
    import torch
    import torch.distributed as dist
    from conftest import load_golden
    from dpsvm_b200 import capi
    os.environ["DPSVM_EXCHANGE"] = exchange
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", init_method="file://" + os.path.join(out_dir, "rdzv"), rank=rank, world_size=world)
    g = load_golden(name)
    n, d = int(g["n"]), int(g["d"])
    size, disp = capi.shard_layout(n, world)
    lo, cnt = int(disp[rank]), int(size[rank])
    ids = [capi.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    with capi.Trainer(n, d, float(g["c"]), float(g["gamma"]), float(g["epsilon"]), max_iter=int(g["max_iter"]), cache_size=cache,
                      shard_rows=cnt, shard_start=lo, device=rank,
                      flags=capi.FLAG_X_SHARD_ONLY | (capi.FLAG_GRAPH_LOOP if graph_loop else 0) | (capi.FLAG_X_MIRROR if mirror else 0)) as t:
        t.comm_init(ids[0], rank, world)
        t.upload(np.ascontiguousarray(g["x"][lo:lo + cnt]), np.ascontiguousarray(g["y"][lo:lo + cnt]))
        st = t.train()
        alpha, f, _ = t.download()
        np.savez(os.path.join(out_dir, "r%d.npz" % rank), alpha=alpha, f=f, it=st.num_iter, conv=st.converged,
                 b=st.b, hits=st.cache_hits, lo=lo, transport=t.comm_transport, resident=t.loop_info()[0])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("name,cache,exchange,layout", [
    ("mnist_512", 0, "p2p", "auto"), ("adult_2048", 10, "p2p", "auto"), ("odd_d_333x37", 4, "p2p", "auto"),
    ("mnist_512", 10, "p2p+graph", "auto"), ("adult_2048", 0, "p2p+graph", "auto"),
    ("mnist_512", 10, "nccl", "auto"), ("odd_d_333x37", 0, "nccl", "auto"),
    # the same exchanges with every rank sweeping the sparse stream of its shard (goldens are small: auto keeps them dense)
    ("mnist_512", 10, "p2p", "sparse"), ("adult_2048", 0, "p2p+graph", "sparse"), ("mnist_512", 0, "nccl", "sparse"),
    # X mirrored on every rank (DPSVM_FLAG_X_MIRROR): 48-byte flagged-word exchange, pair rows from local memory
    ("mnist_512", 0, "p2p-ll", "auto"), ("adult_2048", 10, "p2p-ll", "auto"), ("odd_d_333x37", 4, "p2p-ll", "auto"),
    ("dups_96x16", 2, "p2p-ll", "auto"), ("mnist_512", 10, "p2p-ll+graph", "auto"), ("adult_2048", 0, "p2p-ll+graph", "sparse"),
    ("mnist_1024_tight", 64, "p2p-ll", "sparse")])
def test_sharded_training_equals_single_gpu(golden, tmp_path, monkeypatch, name, cache, exchange, layout):
    import torch.multiprocessing as mp
    from dpsvm_b200 import capi
    monkeypatch.setenv("DPSVM_LAYOUT", layout)
    world = min(_ngpu(), 4 if name == "adult_2048" else (3 if name == "odd_d_333x37" else 2))
    mp.spawn(_worker, args=(world, _free_port(), name, cache, exchange, str(tmp_path), layout), nprocs=world, join=True)
    g = golden(name)
    with capi.Trainer(int(g["n"]), int(g["d"]), float(g["c"]), float(g["gamma"]), float(g["epsilon"]),
                      max_iter=int(g["max_iter"]), cache_size=cache) as t:
        t.upload(np.ascontiguousarray(g["x"]), np.ascontiguousarray(g["y"]))
        st = t.train()
        alpha, f, _ = t.download()
    parts = [np.load(tmp_path / ("r%d.npz" % r)) for r in range(world)]
    for p in parts:
        assert str(p["transport"]) == exchange.split("+")[0]      # peer mapping must actually be in use when asked for
        assert bool(p["resident"]) == (exchange in ("p2p", "p2p-ll"))   # resident loop over peer memory, graph loop otherwise
        assert int(p["it"]) == st.num_iter and int(p["conv"]) == st.converged and float(p["b"]) == st.b
        assert int(p["hits"]) == st.cache_hits
    assert np.array_equal(np.concatenate([p["alpha"] for p in parts]), alpha)
    assert np.array_equal(np.concatenate([p["f"] for p in parts]), f)


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("layout", ["dense", "sparse"])
def test_one_process_two_devices(golden, monkeypatch, layout):
    """The C ABI takes an explicit device ordinal (the reference never calls cudaSetDevice): two contexts on two
    GPUs driven by ONE process give the same result (kernel attributes are set per device, not per process)."""
    from dpsvm_b200 import capi
    monkeypatch.setenv("DPSVM_LAYOUT", layout)
    g = golden("mnist_512")
    out = []
    for dev in (0, 1, 0):
        with capi.Trainer(int(g["n"]), int(g["d"]), float(g["c"]), float(g["gamma"]), float(g["epsilon"]),
                          max_iter=int(g["max_iter"]), cache_size=10, device=dev) as t:
            t.upload(np.ascontiguousarray(g["x"]), np.ascontiguousarray(g["y"]))
            st = t.train()
            alpha, f, _ = t.download()
            out.append((st.num_iter, alpha, f))
    for it, alpha, f in out[1:]:
        assert it == out[0][0] and np.array_equal(alpha, out[0][1]) and np.array_equal(f, out[0][2])
    # wide rows: the generic pass kernel's shared-memory attribute must be set on the second device too
    monkeypatch.setenv("DPSVM_LAYOUT", "auto")        # rows of 1100 floats have no sparse sweep
    rng = np.random.default_rng(1)
    x = rng.random((64, 1100), dtype=np.float32)
    y = np.where(rng.random(64) < 0.5, 1, -1).astype(np.int32)
    its = []
    for dev in (0, 1):
        with capi.Trainer(64, 1100, 1.0, 0.01, 1e-3, cache_size=0, device=dev) as t:
            t.upload(x, y)
            its.append(t.train(50).num_iter)
    assert its[0] == its[1] == 50
