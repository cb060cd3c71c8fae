#!/usr/bin/env python
"""Runs ON a multi-GPU box under torchrun: train a sharded workload for up to --max-iter iterations with
a given kernel-row cache size and report iterations, time, iterations/s and the cache hit rate
(BASELINE config 5: 4M x 128 on 8 GPUs with the cache under HBM pressure).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 \
      tools/converge_multi.py --workload blobs4m --cache 16384 --max-iter 200000
"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="blobs4m")
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--cache", type=int, default=10)
    ap.add_argument("--max-iter", type=int, default=200000)
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    from dpsvm_b200 import capi, datagen

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n_total = args.rows or {"blobs4m": 4000000, "mnist1m": 1000000, "mnist60k": 60000}[args.workload]
    size, disp = capi.shard_layout(n_total, world)
    x, y, prm = datagen.make(args.workload, n=int(size[rank]), row0=int(disp[rank]))
    d = x.shape[1]
    tr = capi.Trainer(n_total, d, prm["c"], prm["gamma"], prm["epsilon"], max_iter=args.max_iter, cache_size=args.cache,
                      shard_rows=int(size[rank]), shard_start=int(disp[rank]), device=local,
                      flags=capi.FLAG_X_SHARD_ONLY if world > 1 else 0)
    if world > 1:
        ids = [capi.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        tr.comm_init(ids[0], rank, world)
    tr.upload(torch.from_numpy(x).pin_memory(), torch.from_numpy(y).pin_memory())
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    st = tr.train()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    alpha, _, _ = tr.download(f=False)
    nsv = torch.tensor([int((alpha > 0).sum())], device="cuda")
    if world > 1:
        dist.all_reduce(nsv)
    resident, chunk = tr.loop_info()
    if rank == 0:
        hit = 100.0 * st.cache_hits / st.cache_lookups if st.cache_lookups else 0.0
        line_mb = 4e-6 * (int(size[rank]) + 31) // 32 * 32
        print("| %s %dx%d | %d | %d lines (%.1f GB per GPU) | %s / %s | %d | %s | %.2f | %.0f | %.1f %% | %d |" % (
            args.workload, n_total, d, world, args.cache, args.cache * 4e-9 * ((int(size[rank]) + 31) // 32 * 32),
            "resident" if resident else "graph", tr.comm_transport, st.num_iter, "yes" if st.converged else "no (max-iter)",
            dt, st.num_iter / dt, hit, int(nsv.item())), flush=True)
    tr.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
