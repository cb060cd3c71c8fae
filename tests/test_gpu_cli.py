"""The binaries, on a GPU: this repo's CLI, the UNMODIFIED reference driver linked against this repo's
SvmTrain (the drop-in proof, tools/dropin_check.sh), and the reference's own seq_test evaluator fed
with a model this repo wrote -- all on a golden data set of the reference's seq.cpp."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT
from dpsvm_b200 import capi, datagen

pytestmark = pytest.mark.gpu

CLI = os.path.join(ROOT, "dpsvm_b200", "svmTrain")
REFMAIN = os.path.join(ROOT, "build", "svmTrain_refmain")
REFGPU = os.path.join(ROOT, "oracle", "_ref", "svmTrain_refgpu")
SEQ_TEST = os.path.join(ROOT, "oracle", "_ref", "seq_test")


def _run(exe, g, csv, model, extra=()):
    cmd = [exe, "-a", str(int(g["d"])), "-x", str(int(g["n"])), "-f", csv, "-m", model, "-c", "%g" % float(g["c"]),
           "-g", "%g" % float(g["gamma"]), "-e", "%g" % float(g["epsilon"]), "-n", str(int(g["max_iter"]))] + list(extra)
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    txt = out.stdout
    it = int(re.search(r"Converged at iteration number: (\d+)", txt).group(1))
    b = float(re.search(r"^b: (\S+)", txt, re.M).group(1))
    nsv = int(re.search(r"Number of SVs: (\d+)", txt).group(1))
    acc = float(re.search(r"Training accuracy: (\S+)", txt).group(1))
    return it, b, nsv, acc, txt


@pytest.fixture()
def case(golden, tmp_path, monkeypatch):
    g = golden("mnist_512")
    monkeypatch.chdir(tmp_path)          # the reference's evaluators keep file names in char[30] / char[60]
    datagen.write_csv("train.csv", g["x"], g["y"])
    datagen.write_csv("test.csv", g["xt"], g["yt"])
    return g


@pytest.mark.skipif(not os.path.exists(CLI), reason="CLI not built")
def test_cli_trains_and_writes_the_reference_model_format(case):
    g = case
    it, b, nsv, acc, txt = _run(CLI, g, "train.csv", "model.txt")
    ref_it = int(g["num_iter"])
    assert abs(it - ref_it) <= max(1, round(0.01 * ref_it))
    assert abs(b - float(g["b"])) < 2e-3 and nsv == int((g["alpha"] != 0).sum()) and acc > 0.99
    for line in ("Populated Data from input file at node: 0", "SETUP DONE", "TOTAL TIME TAKEN in seconds:",
                 "Training model has been saved to the file model.txt"):
        assert line in txt                                     # the stdout lines tools scrape (svmTrainMain.cpp:181-350)
    lines = open("model.txt").read().splitlines()
    assert float(lines[0]) == float(g["gamma"]) and abs(float(lines[1]) - b) < 1e-5     # gamma, b (svmTrainMain.cpp:393-394)
    assert len(lines) - 2 == nsv
    first = lines[2].split(",")
    assert len(first) == 2 + int(g["d"]) and first[1] in ("1", "-1")
    # same run through the library API: the CLI is only a driver
    with capi.Trainer(int(g["n"]), int(g["d"]), float(g["c"]), float(g["gamma"]), float(g["epsilon"]),
                      max_iter=int(g["max_iter"]), cache_size=10) as t:
        t.upload(np.ascontiguousarray(g["x"]), np.ascontiguousarray(g["y"]))
        assert t.train().num_iter == it


@pytest.mark.skipif(not (os.path.exists(CLI) and os.path.exists(SEQ_TEST)), reason="needs the CLI and oracle/_ref/seq_test")
def test_reference_evaluator_accepts_our_model_and_predicts_the_same_labels(case):
    """north_star: held-out predicted labels must match seq.cpp -- checked with the reference's own
    seq_test binary reading a model file written by this repo (--seq-model: gamma line, no b line)."""
    g = case
    _run(CLI, g, "train.csv", "seq.model", extra=["--seq-model"])
    out = subprocess.run([SEQ_TEST, "-a", str(int(g["d"])), "-x", str(int(g["n_test"])), "-f", "test.csv", "-m", "seq.model"],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    acc = float(re.search(r"Test accuracy: (\S+)", out.stdout).group(1))
    want = float((g["test_labels"] == g["yt"]).mean())           # labels seq.cpp's own solution gives on the same rows
    assert abs(acc - want) < 0.5 / int(g["n_test"])


@pytest.mark.skipif(not os.path.exists(REFMAIN), reason="drop-in binary not built (needs /root/reference at build time)")
def test_unmodified_reference_main_runs_on_this_library(case):
    g = case
    it, b, nsv, acc, _ = _run(REFMAIN, g, "train.csv", "model_ref.txt")
    ref_it = int(g["num_iter"])
    assert abs(it - ref_it) <= max(1, round(0.02 * ref_it))     # host eta (sequential sum) vs seq's snrm2: last-bit effects
    assert abs(b - float(g["b"])) < 2e-3 and nsv == int((g["alpha"] != 0).sum()) and acc > 0.99
    if os.path.exists(CLI):
        it2, b2, nsv2, _, _ = _run(CLI, g, "train.csv", "model_hl.txt", extra=["--host-loop"])
        assert (it2, nsv2) == (it, nsv) and abs(b2 - b) < 1e-6   # same loop structure, same library: same run


@pytest.mark.skipif(not os.path.exists(REFGPU), reason="reference GPU path not built")
def test_reference_gpu_path_rebuilt_for_sm100_agrees(case):
    """the reference's own cuBLAS + Thrust code, rebuilt unmodified for sm_100a."""
    g = case
    it, b, nsv, acc, _ = _run(REFGPU, g, "train.csv", "model_rg.txt", extra=["-s", "10"])
    ref_it = int(g["num_iter"])
    assert abs(it - ref_it) <= round(0.03 * ref_it) and abs(b - float(g["b"])) < 2e-3


def test_bench_line_carries_the_contract_and_matches_the_reference_arm(tmp_path):
    """bench.py on a reduced row count (seconds): one JSON line with every key the driver reads, and a `config` object
    identical to the one `--impl reference` prints for the same arguments (the driver compares the two arms' configs)."""
    import json
    import sys
    common = ["--rows", "6000", "--steps", "3", "--warmup", "3"]
    ours = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--e2e-steps", "1", "--converge", "1", "--no-cpu-baseline",
                           "--no-ref-gpu"] + common, capture_output=True, text=True, timeout=600)
    assert ours.returncode == 0, ours.stderr[-3000:]
    d = json.loads([l for l in ours.stdout.splitlines() if l.startswith("{")][-1])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                "dtype", "data", "config", "roofline", "e2e", "time_to_converge", "gpu_launches", "clocks"):
        assert key in d, key
    assert d["metric"] == "smo_iterations_per_s" and d["unit"] == "iter/s" and d["value"] > 0 and d["gpu_launches"] > 0
    for key in ("bound", "achieved", "peak", "unit", "frac", "traffic", "frac_dense_basis"):
        assert key in d["roofline"], key
    assert abs(d["roofline"]["frac"] - d["roofline"]["achieved"] / d["roofline"]["peak"]) < 1e-9
    assert d["e2e"]["h2d_bytes_per_step"] == 4 * 6000 * 784 + 4 * 6000 and d["e2e"]["value"] > 0
    ttc = d["time_to_converge"]
    assert ttc["converged"] is True and ttc["iterations"] > 0 and ttc["n_sv"] > 0 and ttc["seconds"] > 0
    big = d["time_to_converge_hbm_cache"]
    assert big["same_result_as_small_cache"] is True and big["loop"] == "resident" and 0 < big["cache_hit_rate"] < 1
    ref = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference"] + common,
                         capture_output=True, text=True, timeout=600)
    assert ref.returncode == 0, ref.stderr[-2000:]
    r = json.loads([l for l in ref.stdout.splitlines() if l.startswith("{")][-1])
    assert r["config"] == d["config"] and r["metric"] == d["metric"] and r["unit"] == d["unit"]
