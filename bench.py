#!/usr/bin/env python
"""bench.py -- SMO iterations/s on the BASELINE.json workload, one JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]

A "step" is ``iters_per_step`` SMO iterations (one fused pass over the X shard each) launched as
CUDA-graph chunks; ``value`` = iterations/s of the whole job with X resident in HBM; ``e2e`` = the same
metric through the public C-ABI call sequence with HOST buffers (upload of X and y from pinned
memory, the iterations, download of alpha) inside the timed region.

``--impl reference`` times the reference's own sequential CPU implementation (oracle/_ref, built from
/root/reference/seq.cpp; falls back to the oracle port) on the same workload, rank 0 only.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "smo_iterations_per_s"
UNIT = "iter/s"
WORKLOADS = {
    # name: (rows, attrs)  -- BASELINE.json configs[1], [2], [3], [4]
    "mnist60k": (60000, 784),
    "adult32k": (32561, 123),
    "mnist1m": (1000000, 784),
    "blobs4m": (4000000, 128),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="mnist60k", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override the row count (debugging)")
    ap.add_argument("--iters-per-step", type=int, default=256)
    ap.add_argument("--cache-size", type=int, default=10, help="kernel-row cache lines (reference default 10)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-pass-only", action="store_true")
    return ap.parse_args()


def measured_peak():
    """HBM roofline denominator: the driver-measured copy bandwidth, else the profiling guide's fallback."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu_index = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu_index), "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def load_workload(args, rank=0, world=1, shard_only=False):
    from dpsvm_b200 import capi, datagen
    n_total, d = WORKLOADS[args.workload]
    if args.rows:
        n_total = args.rows
    if shard_only and world > 1:
        size, disp = capi.shard_layout(n_total, world)
        x, y, prm = datagen.make(args.workload, n=int(size[rank]), row0=int(disp[rank]))
        return x, y, prm, n_total, d, int(size[rank]), int(disp[rank])
    x, y, prm = datagen.make(args.workload, n=n_total)
    return x, y, prm, n_total, d, n_total, 0


def algorithmic_bytes(n, d, cache_lines, nnz=None):
    """X shard once + |x|^2, f read, f write, alpha, y at 4 B each (logical d); the two
    cache-line stores of a miss/miss iteration add 8 n when the kernel-row cache is on.
    nnz given (sparse stream): X costs 6 B per non-zero (fp32 value + 16-bit column), padding excluded."""
    x_bytes = 4.0 * n * d if nnz is None else 6.0 * nnz
    return x_bytes + 20.0 * n + (8.0 * n if cache_lines >= 2 else 0.0)


def run_reference(args):
    """The reference's own CPU implementation of the path, rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pyoracle
    x, y, prm, n_total, d, _, _ = load_workload(args)
    kind = "reference" if pyoracle.have_ref() else "port"
    total_steps = args.steps + args.warmup
    ips = 4 if total_steps <= 40 else (2 if total_steps <= 100 else 1)      # keeps the whole run within a few minutes

    def step():
        if kind == "reference":
            pyoracle.ref_train(x, y, prm["c"], prm["gamma"], prm["epsilon"], max_iter=ips)
        else:
            pyoracle.train(x, y, prm["c"], prm["gamma"], prm["epsilon"], max_iter=ips, kernel_form=0, select_rule=0)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = args.steps * ips / dt
    sample = ("%d steps x %d SMO iterations of seq.cpp's loop from alpha=0 on the full %dx%d set (setup copy included, "
              "its O(N nSV d) accuracy pass excluded)" % (args.steps, ips, n_total, d))
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "%s %dx%d RBF c=%g gamma=%g eps=%g" % (args.workload, n_total, d, prm["c"], prm["gamma"], prm["epsilon"]),
                   "iters_per_step": ips},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": kind, "sample": sample,
                         "host_cores_available": os.cpu_count()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


def cpu_baseline(x, y, prm, budget_iters=16):
    """Bounded CPU sample on the same workload (about 10-30 s): the reference's seq.cpp loop."""
    from oracle import pyoracle
    kind = "reference" if pyoracle.have_ref() else "port"
    fn = pyoracle.ref_train if kind == "reference" else pyoracle.train
    t0 = time.perf_counter()
    fn(x, y, prm["c"], prm["gamma"], prm["epsilon"], max_iter=1)
    t1 = time.perf_counter() - t0
    iters = max(2, min(budget_iters, int(12.0 / max(t1, 1e-3))))
    t0 = time.perf_counter()
    fn(x, y, prm["c"], prm["gamma"], prm["epsilon"], max_iter=iters)
    dt = time.perf_counter() - t0
    return {"value": iters / dt, "unit": UNIT, "cores": 1, "kind": kind, "host_cores_available": os.cpu_count(),
            "sample": "%d SMO iterations of seq.cpp's loop from alpha=0 on the full %dx%d set, %.1f s wall "
                      "(single-threaded program; 1 of %d host cores)" % (iters, x.shape[0], x.shape[1], dt, os.cpu_count())}


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist
    from dpsvm_b200 import capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    x, y, prm, n_total, d, n_shard, row0 = load_workload(args, rank, world, shard_only=True)
    ips = args.iters_per_step
    max_iter = 2 ** 31 - 2
    flags = capi.FLAG_X_SHARD_ONLY if world > 1 else 0
    tr = capi.Trainer(n_total, d, prm["c"], prm["gamma"], prm["epsilon"], max_iter=max_iter, cache_size=args.cache_size,
                      shard_rows=n_shard, shard_start=row0, device=local_rank, flags=flags)
    hx = torch.from_numpy(x).pin_memory()
    hy = torch.from_numpy(y).pin_memory()
    if world > 1:
        ids = [capi.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        tr.comm_init(ids[0], rank, world)
    tr.upload(hx, hy)

    stream = torch.cuda.ExternalStream(tr.stream)
    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident arm -----------------------------------------------------------------
    for _ in range(args.warmup):
        tr.train_async(ips)
    st_w = tr.sync()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    barrier()
    ev0.record(stream)
    for _ in range(args.steps):
        tr.train_async(ips)
    ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    clk = clocks.stop() if rank == 0 else None
    st = tr.sync()
    done_iters = st.num_iter - st_w.num_iter
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = done_iters / (ms * 1e-3)

    # ---- the dominant kernel alone (stateless launcher, back to back on one stream) -----------
    pass_ms = None
    lanes, nnz, sweep_bytes = tr.layout_info()
    if not args.no_pass_only and lanes == 0:        # the stateless pass entry point sweeps dense rows
        s = tr.device_state()
        pitch = s["pitch"]
        keys = torch.full((2,), -1, dtype=torch.int64, device="cuda")
        xsq = torch.empty(n_shard, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        # the f-vector is scratch here (delta = 0 leaves the values unchanged; the traffic is the same)
        ftmp = torch.zeros(n_shard, dtype=torch.float32, device="cuda")
        lines = torch.empty(2, n_shard, dtype=torch.float32, device="cuda") if args.cache_size >= 2 else None
        capi.row_norms(s["x"], n_shard, d, pitch, xsq, tr.stream)
        tr.sync()

        def one_pass(i):
            # same kernel, same L2 policy and sweep order as iteration i of the trainer
            r_hi, r_lo = (7 * i + 1) % n_shard, (13 * i + 5) % n_shard
            capi.smo_pass(stream=torch.cuda.current_stream().cuda_stream, d_x=s["x"], n=n_shard, d=d, pitch=pitch, d_xsq=xsq,
                          d_y=s["y"], d_alpha=s["alpha"], d_f=ftmp, d_x_hi=s["x"] + 4 * pitch * r_hi,
                          d_x_lo=s["x"] + 4 * pitch * r_lo, xsq_hi=1.0, xsq_lo=1.0, delta_hi=0.0, delta_lo=0.0,
                          gamma=prm["gamma"], c=prm["c"], index_base=row0, d_keys=keys,
                          d_line_hi=lines[0] if lines is not None else None, d_line_lo=lines[1] if lines is not None else None,
                          variant=capi.pass_variant(n_shard, pitch, i))
        # replayed from a CUDA graph so the host's per-launch cost cannot set the pace
        reps = 200
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            one_pass(0)
            one_pass(1)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(reps):
                one_pass(i)
        g.replay()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        pass_ms = e0.elapsed_time(e1) / reps

    # ---- end to end: host buffers in, host buffers out, every step ------------------------------
    e2e_iters = max(ips, 1024)          # a training run is tens of thousands of iterations per upload; 1024 is conservative
    e2e_t = []
    for k in range(args.e2e_steps + 1):
        barrier()
        t0 = time.perf_counter()
        tr.upload(hx, hy)                 # H2D of X and y from pinned host memory, |x|^2, first selection
        tr.train_async(e2e_iters)
        alpha, _, _ = tr.download(alpha=True, f=False)    # D2H of alpha (synchronises)
        st_e = tr.sync()
        barrier()
        if k > 0:                          # first one is a warm-up
            e2e_t.append(time.perf_counter() - t0)
        assert st_e.num_iter == e2e_iters or st_e.converged
    e2e_s = max(e2e_t)
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = e2e_iters / e2e_s

    if rank != 0:
        tr.close()
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    abytes = algorithmic_bytes(n_shard, d, args.cache_size, nnz if lanes else None)
    dense_abytes = algorithmic_bytes(n_shard, d, args.cache_size)
    iter_ms = ms / max(done_iters, 1)
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            key = "%s%s_per_iteration" % (args.workload, "_sparse" if lanes else "")
            per_iter = json.load(fh).get(key if world == 1 else "%s%s_p%d_per_iteration" % (args.workload, "_sparse" if lanes else "", world))
            traffic = per_iter
    except Exception:
        pass
    resident, chunk = tr.loop_info()
    if resident:
        # the dominant kernel IS the loop: one smo_resident_kernel launch = `chunk` whole iterations
        launches = args.steps * ((ips + chunk - 1) // chunk)
        kernel_name = "smo_resident_kernel"
        kernel_ms = ms / launches
        bytes_per_launch = abytes * done_iters / launches
        timing = ("the timed region itself: %d launches of %d iterations each, CUDA events on the launching stream "
                  "(max over ranks)" % (launches, chunk))
        share = 1.0
        gpu_launches = launches
    else:
        kernels_per_iter = 3 if tr.comm_transport == "nccl" else 2     # control + pass (+ pack before the NCCL all-gather)
        kernel_name = "smo_pass_sparse_kernel" if lanes else "smo_pass_kernel"
        kernel_ms = pass_ms if pass_ms is not None else iter_ms
        bytes_per_launch = abytes
        timing = ("pass kernel alone: 200 launches replayed from one CUDA graph with the trainer's L2 policy and sweep order, "
                  "CUDA events on the replay stream") if pass_ms is not None else "whole iteration (control + pass) from the timed region"
        share = (kernel_ms / iter_ms) if pass_ms is not None else 1.0
        gpu_launches = int(done_iters * kernels_per_iter)
    achieved = bytes_per_launch / (kernel_ms * 1e-3) / 1e9
    if traffic is not None:
        traffic = traffic * (bytes_per_launch / abytes)      # ncu DRAM bytes per iteration x iterations per launch
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": "%s %dx%d RBF c=%g gamma=%g eps=%g" % (args.workload, n_total, d, prm["c"], prm["gamma"], prm["epsilon"]),
            "iters_per_step": ips, "iterations_timed": done_iters, "converged_during_timing": bool(st.converged),
            "cache_lines": args.cache_size, "cache_hits": int(st.cache_hits), "cache_lookups": int(st.cache_lookups),
            "rows_per_gpu": n_shard, "us_per_iteration": 1e3 * iter_ms, "exchange": tr.comm_transport,
            "loop": "resident (persistent cooperative kernel, %d iterations per launch)" % chunk if resident
                    else "CUDA graph of [control, pass] x %d" % chunk,
            "pass_kernel_alone_us": (1e3 * pass_ms) if pass_ms is not None else None,
            "layout": ("sparse stream, %d lanes per row: %.1f %% of the cells are non-zero, %.1f MB per sweep (padding included) instead of %.1f MB of dense rows"
                       % (lanes, 100.0 * nnz / (n_shard * d), 1e-6 * sweep_bytes, 4e-6 * n_shard * d)) if lanes
                      else "dense rows, %.1f MB per sweep" % (1e-6 * sweep_bytes),
            "l2": "%.1f MB per sweep vs 126 MB L2: %s; consecutive iterations sweep in opposite directions, nothing is flushed between them (the loop re-reads the same X by construction)"
                  % (1e-6 * sweep_bytes, "inputs larger than L2" if sweep_bytes > 126e6 else "L2-resident"),
            "iteration_GBps_incl_control": abytes / (iter_ms * 1e-3) / 1e9,
        },
        "roofline": {"bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": bytes_per_launch, "algorithmic_bytes_per_iteration": abytes,
                     "algorithmic_bytes_basis": ("sparse stream: 6 B per non-zero + 20 B per row (+ 8 B per row of cache-line stores)" if lanes
                                                 else "SURVEY 8(d): 4 n d + 20 n (+ 8 n of cache-line stores)"),
                     "dense_rows_equivalent_GBps": dense_abytes / (iter_ms * 1e-3) / 1e9,
                     "note": ("the sparse sweep is limited by its shared-memory gathers (LSU), and a stream of up to ~120 MB is served from L2 "
                              "and shared memory, not HBM (profiles/r01_sparse_ncu_summary.md); frac is still quoted against the HBM copy peak") if lanes else None,
                     "kernel_us": 1e3 * kernel_ms, "timing": timing, "share_of_step": share},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(4 * n_shard * d + 4 * n_shard),
                "d2h_bytes_per_step": int(4 * n_shard + 64), "iters_per_step": e2e_iters,
                "what": "dpsvm_upload (X,y from pinned host) + dpsvm_train_async + dpsvm_download(alpha) + dpsvm_sync, wall clock, worst of %d" % args.e2e_steps},
        "gpu_launches": gpu_launches,
        "clocks": clk,
    }
    if not args.no_cpu_baseline and world == 1:
        out["cpu_baseline"] = cpu_baseline(x, y, prm)
    print(json.dumps(out), flush=True)
    tr.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
