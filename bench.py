#!/usr/bin/env python
"""bench.py -- SMO iterations/s and time-to-converge on the BASELINE.json workload, one JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]

A "step" of the device-timed arm is ``iters_per_step`` SMO iterations (one fused pass over the X shard each);
``value`` = iterations/s of the whole job with X resident in HBM.  ``e2e`` = the same metric through the public
C-ABI call sequence with HOST buffers inside the timed region: upload of X and y from pinned memory, a whole
training run (to convergence on the workloads that converge within a second or two, else a fixed number of
iterations), download of alpha.  ``time_to_converge`` comes from those same runs.

With N > 1 ranks the line also carries ``parity`` (sharded run against a single-GPU run of the same iterations
on rank 0, bit for bit, with the layout pinned) and ``n1`` / ``speedup_vs_n1`` (the same workload timed on rank 0
alone in the same process).

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
import zlib

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
CONVERGES_FAST = ("mnist60k", "adult32k")      # a whole training run is a second or two: e2e and time-to-converge train to the end


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
    ap.add_argument("--converge", type=int, default=-1, help="1/0: train to convergence in the e2e runs (default: only the workloads that converge in seconds)")
    ap.add_argument("--parity-iters", type=int, default=2000, help="iterations of the N>1 bit-identity check (0 = skip)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-pass-only", action="store_true")
    ap.add_argument("--no-ref-gpu", action="store_true")
    ap.add_argument("--no-mirror", action="store_true", help="N>1: keep only the owned rows of X on each GPU (candidate rows then travel in the exchange)")
    ap.add_argument("--no-n1", action="store_true", help="N>1: skip the single-GPU run of the same workload on rank 0")
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


def workload_config(args, n_total, d, prm):
    """The `config` object: what is being solved.  Identical in both arms (ours / --impl reference); everything
    that describes HOW a run went is under `run`."""
    return {
        "workload": "%s %dx%d RBF c=%g gamma=%g eps=%g" % (args.workload, n_total, d, prm["c"], prm["gamma"], prm["epsilon"]),
        "cache_lines": args.cache_size,
        "l2": "nothing is flushed between iterations: every SMO iteration re-reads the same X by construction; whether one "
              "sweep exceeds the GPU's 126 MB L2 is stated under run.l2 (GPU arm)",
    }


def algorithmic_bytes(n, d, cache_lines, nnz=None, sweep_bytes=0):
    """X shard once + |x|^2, f read, f write, alpha, y at 4 B each (logical d); the two
    cache-line stores of a miss/miss iteration add 8 n when the kernel-row cache is on.
    nnz given (sparse stream): X costs 5 or 6 B per non-zero (fp32 value + 8-bit column distance or 16-bit column
    offset), padding excluded.  A padded stream of fewer than 6 B per non-zero can only be the 8-bit format."""
    per_nz = 5.0 if (nnz and 0 < sweep_bytes < 6.0 * nnz) else 6.0
    x_bytes = 4.0 * n * d if nnz is None else per_nz * nnz
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
        "config": workload_config(args, n_total, d, prm),
        "run": {"iters_per_step": ips, "rows": n_total, "program": "seq.cpp (sequential SMO, no kernel-row cache: -s does not apply)"},
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


def ref_gpu_baseline(args, x, y, prm, cap=4000):
    """the reference's OWN GPU path (svmTrain.cu + cache.cu: cuBLAS gemv + Thrust), rebuilt unmodified for
    sm_100a as oracle/_ref/svmTrain_refgpu, on the same rows through a CSV file; loop time between its "SETUP DONE" and
    "TOTAL TIME TAKEN" lines, first `cap` iterations."""
    exe = os.path.join(ROOT, "oracle", "_ref", "svmTrain_refgpu")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/svmTrain_refgpu not built (make refgpu needs /root/reference)"}
    from dpsvm_b200 import datagen
    import tempfile
    tmp = tempfile.mkdtemp(prefix="dpsvm_refgpu_")
    csv = os.path.join(tmp, "train.csv")
    try:
        datagen.write_csv(csv, x, y)
        cmd = ["stdbuf", "-oL", exe, "-a", str(x.shape[1]), "-x", str(x.shape[0]), "-f", csv, "-c", "%g" % prm["c"], "-g", "%g" % prm["gamma"],
               "-e", "%g" % prm["epsilon"], "-n", str(cap), "-s", str(max(args.cache_size, 2)), "-m", os.path.join(tmp, "model.txt")]
        p = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, bufsize=1)
        t_setup = t_total = iters = None
        for line in p.stdout:
            now = time.perf_counter()
            if line.startswith("SETUP DONE"):
                t_setup = now
            elif line.startswith("TOTAL TIME TAKEN"):
                t_total = now
            elif "iteration number" in line or "Could not converge in" in line:
                iters = int([w for w in line.replace(".", " ").split() if w.isdigit()][0])
        p.wait(timeout=600)
        if p.returncode != 0 or not (t_setup and t_total and iters):
            return {"unavailable": "svmTrain_refgpu exited with %s" % p.returncode}
        secs = t_total - t_setup
        return {"value": iters / secs, "unit": UNIT, "iterations": iters, "loop_seconds": secs,
                "what": "the reference's own GPU path (svmTrain.cu + cache.cu, cuBLAS sgemv + Thrust reductions, host loop), compiled "
                        "unmodified for sm_100a, same rows via CSV, first %d iterations, wall clock between its SETUP DONE and TOTAL TIME "
                        "TAKEN lines (includes its final accuracy print)" % cap}
    except Exception as e:      # a comparison leg must never cost the bench line
        return {"unavailable": "%s: %s" % (type(e).__name__, e)}
    finally:
        for f in ("train.csv", "model.txt"):
            try:
                os.remove(os.path.join(tmp, f))
            except OSError:
                pass
        try:
            os.rmdir(tmp)
        except OSError:
            pass


def _crc(a):
    return zlib.crc32(np.ascontiguousarray(a).view(np.uint8))


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

    # each rank uploads only its rows; the ranks then mirror one another's rows device to device when that fits easily
    # (every BASELINE config does: at most 3.1 GB), so the selected pair's rows are local on every GPU
    mirror = world > 1 and not args.no_mirror and 4.0 * n_total * capi.pitch_for(d) <= 40e9

    def make_trainer(extra_flags=0, single=False, cache_lines=None):
        """A trainer for this rank's shard (joined to the other ranks), or -- single -- for the whole set on this GPU."""
        cache_lines = args.cache_size if cache_lines is None else cache_lines
        if single or world == 1:
            return capi.Trainer(n_total, d, prm["c"], prm["gamma"], prm["epsilon"], max_iter=max_iter, cache_size=cache_lines,
                                device=local_rank, flags=extra_flags)
        t = capi.Trainer(n_total, d, prm["c"], prm["gamma"], prm["epsilon"], max_iter=max_iter, cache_size=cache_lines,
                         shard_rows=n_shard, shard_start=row0, device=local_rank,
                         flags=capi.FLAG_X_SHARD_ONLY | (capi.FLAG_X_MIRROR if mirror else 0) | extra_flags)
        ids = [capi.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        t.comm_init(ids[0], rank, world)
        return t

    tr = make_trainer()
    hx = torch.from_numpy(x).pin_memory()
    hy = torch.from_numpy(y).pin_memory()
    tr.upload(hx, hy)

    stream = torch.cuda.ExternalStream(tr.stream)
    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], device="cuda", dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return int(t.item())

    def timed_iterations(t, steps, warmup, sample_clocks=False):
        """`steps` x ips iterations on the trainer's stream between CUDA events, after `warmup` untimed steps."""
        s = torch.cuda.ExternalStream(t.stream)
        for _ in range(warmup):
            t.train_async(ips)
        st_w = t.sync()
        clocks = ClockSampler(local_rank) if sample_clocks else None
        if clocks:
            clocks.start()
        barrier()
        ev0.record(s)
        for _ in range(steps):
            t.train_async(ips)
        ev1.record(s)
        barrier()
        ms = ev0.elapsed_time(ev1)
        clk = clocks.stop() if clocks else None
        st = t.sync()
        return ms, st.num_iter - st_w.num_iter, st, clk

    # ---- device-resident arm -----------------------------------------------------------------
    ms, done_iters, st, clk = timed_iterations(tr, args.steps, args.warmup, sample_clocks=(rank == 0))
    ms = max_over_ranks(ms)
    value = done_iters / (ms * 1e-3)

    # ---- the dominant kernel alone (stateless launcher, back to back on one stream) -----------
    pass_ms = None
    lanes, nnz, sweep_bytes = tr.layout_info()
    if not args.no_pass_only and lanes == 0 and world == 1:        # the stateless pass entry point sweeps dense rows
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
    # One e2e step is one whole job as a user runs it: dpsvm_upload -> dpsvm_train -> dpsvm_download.
    to_convergence = (args.workload in CONVERGES_FAST and not args.rows) if args.converge < 0 else bool(args.converge)
    e2e_cap = 0 if to_convergence else max(ips, 1024)
    e2e_runs = []
    for k in range(args.e2e_steps + 1):
        barrier()
        t0 = time.perf_counter()
        tr.upload(hx, hy)                 # H2D of X and y from pinned host memory, |x|^2, sparse stream, first selection
        t1 = time.perf_counter()
        st_e = tr.train(e2e_cap)          # the loop, to convergence (or e2e_cap iterations)
        t2 = time.perf_counter()
        alpha, _, _ = tr.download(alpha=True, f=False)    # D2H of alpha (synchronises)
        barrier()
        t3 = time.perf_counter()
        if k > 0:                          # first one is a warm-up
            e2e_runs.append((t3 - t0, t2 - t1, t1 - t0, st_e.num_iter, st_e.converged, st_e.b))
        assert st_e.converged or st_e.num_iter == e2e_cap
    e2e_runs.sort()
    e2e_s, train_s, upload_s, e2e_iters, e2e_conv, e2e_b = e2e_runs[-1]        # the worst run
    e2e_s = max_over_ranks(e2e_s)
    train_s = max_over_ranks(train_s)
    e2e_value = e2e_iters / e2e_s
    n_sv = sum_over_ranks(int((alpha > 0).sum()))
    n_bound = sum_over_ranks(int((alpha == np.float32(prm["c"])).sum()))
    st_last = tr.sync()
    time_to_converge = {
        "seconds": train_s, "seconds_with_upload_and_download": e2e_s, "upload_and_setup_seconds": max_over_ranks(upload_s),
        "iterations": int(e2e_iters), "converged": bool(e2e_conv), "n_sv": n_sv, "n_sv_at_c": n_bound, "b": float(e2e_b),
        "cache_lines": args.cache_size, "cache_hits": int(st_last.cache_hits), "cache_lookups": int(st_last.cache_lookups),
        "what": ("one dpsvm_train call from alpha = 0 to b_lo <= b_hi + 2 eps (svmTrainMain.cpp:235-310), X resident; wall clock, "
                 "worst of %d runs, max over ranks" % args.e2e_steps) if to_convergence else
                ("NOT run to convergence on this workload by default (--converge 1 does): first %d iterations" % e2e_cap),
    }

    # ---- the same job with the kernel-row cache sized to HBM (one line per row: nothing is ever evicted) ----
    big_cache = None
    if to_convergence and 4.0 * n_total * n_shard <= 40e9:
        with make_trainer(cache_lines=n_total) as tb:
            runs = []
            for k in range(2):
                barrier()
                tb.upload(hx, hy)
                t1 = time.perf_counter()
                sb = tb.train(0)
                t2 = time.perf_counter()
                runs.append(t2 - t1)
            a_b, _, _ = tb.download(alpha=True, f=False)
            big_cache = {"seconds": max_over_ranks(min(runs)), "iterations": int(sb.num_iter), "converged": bool(sb.converged),
                         "cache_lines": n_total, "cache_GB_per_gpu": 4e-9 * n_total * n_shard,
                         "cache_hit_rate": float(sb.cache_hits) / max(1, int(sb.cache_lookups)), "b": float(sb.b),
                         "same_result_as_small_cache": bool(int(sb.num_iter) == int(e2e_iters) and float(sb.b) == float(e2e_b)
                                                            and np.array_equal(a_b, alpha)),
                         "loop": "resident" if tb.loop_info()[0] else "graph",
                         "what": "dpsvm_train to convergence with -s N (a cache line for every row, %d x %d floats per GPU): an iteration whose two "
                                 "kernel rows are both cached reads 28 B per row instead of X; second of two runs" % (n_total, n_shard)}

    # ---- N > 1: the same workload on ONE GPU in the same process (rank 0), and bit-identity of the sharded run ----
    parity = None
    n1 = None
    if world > 1 and not (args.no_n1 and args.parity_iters <= 0):
        pitch = capi.pitch_for(d)
        size, disp = capi.shard_layout(n_total, world)
        # rank 0 collects every shard over NCCL into one device buffer (no second host copy of the whole set)
        mine = torch.zeros(n_shard, pitch, dtype=torch.float32, device="cuda")
        mine[:, :d] = hx.cuda()
        ymine = hy.cuda()
        full_x = full_y = None
        if rank == 0:
            full_x = torch.empty(n_total, pitch, dtype=torch.float32, device="cuda")
            full_y = torch.empty(n_total, dtype=torch.int32, device="cuda")
            full_x[:n_shard] = mine
            full_y[:n_shard] = ymine
            for r in range(1, world):
                dist.recv(full_x[int(disp[r]):int(disp[r]) + int(size[r])], src=r)
                dist.recv(full_y[int(disp[r]):int(disp[r]) + int(size[r])], src=r)
        else:
            dist.send(mine, dst=0)
            dist.send(ymine, dst=0)
        del mine
        barrier()

        if not args.no_n1 and rank == 0:
            with make_trainer(single=True) as t1:
                t1.upload_device(full_x, pitch, full_y)
                steps1 = max(3, args.steps // 2)
                ms1, it1, _, _ = timed_iterations_single(torch, t1, ips, steps1, 3)
                l1, _, sb1 = t1.layout_info()
                n1 = {"value": it1 / (ms1 * 1e-3), "unit": UNIT, "us_per_iteration": 1e3 * ms1 / max(it1, 1), "steps": steps1,
                      "layout": ("sparse stream, %d lanes" % l1) if l1 else "dense rows", "sweep_MB": 1e-6 * sb1,
                      "what": "the same workload on rank 0's GPU alone, same process, X resident, CUDA events"}
        barrier()

        if args.parity_iters > 0:
            k_par = args.parity_iters
            dense_frac = float(np.count_nonzero(x)) / max(1, x.size)
            pins = [("dense rows", capi.FLAG_DENSE, None)]
            if dense_frac < 0.7 and d <= 1024:
                pins.append(("sparse stream, 8 lanes per row", capi.FLAG_SPARSE, "8"))
            checks = []
            for name, flag, lanes_env in pins:
                old = os.environ.get("DPSVM_SPARSE_LANES")
                if lanes_env:
                    os.environ["DPSVM_SPARSE_LANES"] = lanes_env
                with make_trainer(extra_flags=flag) as tp:
                    tp.upload(hx, hy)
                    sp = tp.train(k_par)
                    a_p, f_p, _ = tp.download()
                    rec = (int(sp.num_iter), float(sp.b_hi), float(sp.b_lo), int(sp.i_hi), int(sp.i_lo), _crc(a_p), _crc(f_p),
                           row0, n_shard, tp.comm_transport, bool(tp.loop_info()[0]))
                recs = [None] * world
                dist.all_gather_object(recs, rec)
                ok = None
                if rank == 0:
                    with make_trainer(extra_flags=flag, single=True) as ts:
                        ts.upload_device(full_x, pitch, full_y)
                        ss = ts.train(k_par)
                        a_s, f_s, _ = ts.download()
                    ok = True
                    for r in recs:
                        lo, cnt = r[7], r[8]
                        ok = ok and r[:5] == (int(ss.num_iter), float(ss.b_hi), float(ss.b_lo), int(ss.i_hi), int(ss.i_lo))
                        ok = ok and r[5] == _crc(a_s[lo:lo + cnt]) and r[6] == _crc(f_s[lo:lo + cnt])
                    checks.append({"layout": name, "bit_identical": bool(ok), "iterations": int(ss.num_iter),
                                   "pair_of_last_iteration": [int(ss.i_hi), int(ss.i_lo)], "exchange": recs[0][9],
                                   "loop": "resident" if recs[0][10] else "graph"})
                if lanes_env:
                    if old is None:
                        os.environ.pop("DPSVM_SPARSE_LANES", None)
                    else:
                        os.environ["DPSVM_SPARSE_LANES"] = old
                barrier()
            if rank == 0:
                parity = {"checked": True, "bit_identical": all(c["bit_identical"] for c in checks), "k": k_par, "runs": checks,
                          "what": "every rank trains k iterations on its shard with the layout pinned; rank 0 repeats the same k iterations "
                                  "on ONE GPU holding the whole set; per rank (num_iter, b_hi, b_lo, i_hi, i_lo) are compared and the CRC-32 "
                                  "of the alpha and f shards against the same slices of the single-GPU arrays"}
        del full_x, full_y

    if rank != 0:
        tr.close()
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    stream_abytes = algorithmic_bytes(n_shard, d, args.cache_size, nnz if lanes else None, sweep_bytes)
    dense_abytes = algorithmic_bytes(n_shard, d, args.cache_size)
    iter_ms = ms / max(done_iters, 1)
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            tj = json.load(fh)
        key = "%s%s%s" % (args.workload, "_sparse" if lanes else "", "" if world == 1 else "_p%d" % world)
        ent = tj.get("entries", {}).get(key)
        if ent:
            traffic, traffic_src = float(ent["dram_bytes_per_iteration"]), {k: ent.get(k) for k in ("commit", "source")}
    except Exception:
        pass
    resident, chunk = tr.loop_info()
    if resident:
        # the dominant kernel IS the loop: one smo_resident_kernel launch = `chunk` whole iterations
        launches = args.steps * ((ips + chunk - 1) // chunk)
        kernel_name = "smo_resident_kernel"
        kernel_ms = ms / launches
        iters_per_launch = done_iters / launches
        timing = ("the timed region itself: %d launches of %d iterations each, CUDA events on the launching stream "
                  "(max over ranks)" % (launches, chunk))
        share = 1.0
        gpu_launches = launches
    else:
        kernels_per_iter = 3 if tr.comm_transport == "nccl" else 2     # control + pass (+ pack before the NCCL all-gather)
        kernel_name = "smo_pass_sparse_kernel" if lanes else "smo_pass_kernel"
        kernel_ms = pass_ms if pass_ms is not None else iter_ms
        iters_per_launch = 1.0
        timing = ("pass kernel alone: 200 launches replayed from one CUDA graph with the trainer's L2 policy and sweep order, "
                  "CUDA events on the replay stream") if pass_ms is not None else "whole iteration (control + pass) from the timed region"
        share = (kernel_ms / iter_ms) if pass_ms is not None else 1.0
        gpu_launches = int(done_iters * kernels_per_iter)
    # SURVEY 8(d) basis (dense rows, 4 n d + 20 n): the figure `achieved` is defined on.  When the sweep reads the sparse
    # stream instead, the same time also gives the stream's own bytes (6 B per non-zero): both fractions are reported.
    achieved = dense_abytes * iters_per_launch / (kernel_ms * 1e-3) / 1e9
    achieved_stream = stream_abytes * iters_per_launch / (kernel_ms * 1e-3) / 1e9
    on_chip = (not lanes) and sweep_bytes <= 30e6
    if lanes:
        bound = "l2+lsu" if sweep_bytes <= 126e6 else "hbm"
    else:
        bound = "smem" if on_chip else "hbm"
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, n_total, d, prm),
        "run": {
            "iters_per_step": ips, "iterations_timed": done_iters, "converged_during_timing": bool(st.converged),
            "cache_hits": int(st.cache_hits), "cache_lookups": int(st.cache_lookups),
            "rows_per_gpu": n_shard, "us_per_iteration": 1e3 * iter_ms, "exchange": tr.comm_transport,
            "x_residency": ("each GPU sweeps its own %d rows and holds a read-only mirror of the other ranks' rows (%.2f GB in all) for the pair fetch"
                            % (n_shard, 4e-9 * n_total * capi.pitch_for(d))) if mirror else ("owned rows only" if world > 1 else "whole set"),
            "loop": "resident (persistent cooperative kernel, %d iterations per launch)" % chunk if resident
                    else "CUDA graph of [control, pass] x %d" % chunk,
            "pass_kernel_alone_us": (1e3 * pass_ms) if pass_ms is not None else None,
            "layout": ("sparse stream, %d lanes per row: %.1f %% of the cells are non-zero, %.1f MB per sweep (padding included) instead of %.1f MB of dense rows"
                       % (lanes, 100.0 * nnz / (n_shard * d), 1e-6 * sweep_bytes, 4e-6 * n_shard * d)) if lanes
                      else "dense rows, %.1f MB per sweep" % (1e-6 * sweep_bytes),
            "l2": "%.1f MB per sweep vs 126 MB L2: %s; consecutive iterations sweep in opposite directions"
                  % (1e-6 * sweep_bytes, "inputs larger than L2" if sweep_bytes > 126e6 else "L2-resident"),
        },
        "roofline": {"bound": bound, "kernel": kernel_name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": (traffic * iters_per_launch) if traffic is not None else None,
                     "traffic_source": traffic_src, "peak_source": peak_src,
                     "algorithmic_bytes_basis": "SURVEY 8(d): 4 n d + 20 n per iteration (+ 8 n of cache-line stores), dense rows, logical d",
                     "algorithmic_bytes_per_iteration": dense_abytes, "algorithmic_bytes_per_launch": dense_abytes * iters_per_launch,
                     "frac_dense_basis": achieved / peak,
                     "frac_stream_basis": (achieved_stream / peak) if lanes else None,
                     "stream_bytes_per_iteration": stream_abytes if lanes else None,
                     "note": ("the sweep reads the sparse stream (5 or 6 B per non-zero, zeros skipped), so on the dense-row basis the figure can exceed "
                              "the HBM peak; a stream of up to ~120 MB is served from L2 and the busiest units are the "
                              "shared-memory gathers (LSU) and the L2->SM fabric, not HBM -- `bound` says which; frac_stream_basis counts only "
                              "the bytes the kernel is asked to read") if lanes else
                             ("the whole shard is held in shared memory across iterations: nothing streams from HBM" if on_chip else None),
                     "kernel_us": 1e3 * kernel_ms, "iterations_per_launch": iters_per_launch, "timing": timing, "share_of_step": share},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(4 * n_shard * d + 4 * n_shard),
                "d2h_bytes_per_step": int(4 * n_shard + 2 * 128), "iters_per_step": int(e2e_iters), "seconds_per_step": e2e_s,
                "what": "one e2e step = one whole job through the C ABI with host buffers: dpsvm_upload (X, y from pinned host memory; |x|^2, "
                        "sparse stream, first selection) + dpsvm_train (%s) + dpsvm_download(alpha); wall clock, worst of %d, max over ranks.  "
                        "Not the 256-iteration step of `value`: a job uploads X once and then iterates"
                        % ("to convergence" if to_convergence else "%d iterations" % e2e_cap, args.e2e_steps)},
        "time_to_converge": time_to_converge,
        "time_to_converge_hbm_cache": big_cache,
        "gpu_launches": gpu_launches,
        "clocks": clk,
    }
    if parity is not None:
        out["parity"] = parity
    if n1 is not None:
        out["n1"] = n1
        out["speedup_vs_n1"] = value / n1["value"]
    if world == 1 and not args.no_ref_gpu and not args.rows and args.workload in CONVERGES_FAST:
        tr.close()
        tr = None
        out["ref_gpu_baseline"] = ref_gpu_baseline(args, x, y, prm)
    if not args.no_cpu_baseline and world == 1:
        out["cpu_baseline"] = cpu_baseline(x, y, prm)
    print(json.dumps(out), flush=True)
    if tr is not None:
        tr.close()
    if world > 1:
        dist.destroy_process_group()


def timed_iterations_single(torch, t, ips, steps, warmup):
    """rank 0 alone (no collective inside): `steps` x ips iterations between CUDA events on the trainer's stream."""
    s = torch.cuda.ExternalStream(t.stream)
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    for _ in range(warmup):
        t.train_async(ips)
    st_w = t.sync()
    torch.cuda.synchronize()
    e0.record(s)
    for _ in range(steps):
        t.train_async(ips)
    e1.record(s)
    torch.cuda.synchronize()
    st = t.sync()
    return e0.elapsed_time(e1), st.num_iter - st_w.num_iter, st, None


if __name__ == "__main__":
    main()
