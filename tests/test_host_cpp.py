"""The C++ host layer without a GPU: populate_data (reference parse.cpp:10-43) and the model files
(reference svmTrainMain.cpp:386-416 / seq.cpp:295-321), through tests/cpp/test_host.cpp linked against
libdpsvm_host.so.  The model writer is compared byte for byte with the file the UNMODIFIED reference
`seq` binary writes for the same alphas."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT
from dpsvm_b200 import datagen

EXE = os.path.join(ROOT, "build", "test_host")
SEQ = os.path.join(ROOT, "oracle", "_ref", "seq")


@pytest.fixture(scope="module")
def exe():
    os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
    if not os.path.exists(os.path.join(ROOT, "dpsvm_b200", "libdpsvm_host.so")):
        subprocess.run(["make", "-C", ROOT, "host"], check=True, capture_output=True)
    subprocess.run(["g++", "-O1", "-std=c++17", "-I" + os.path.join(ROOT, "include"), "-o", EXE,
                    os.path.join(ROOT, "tests", "cpp", "test_host.cpp"), "-L" + os.path.join(ROOT, "dpsvm_b200"),
                    "-ldpsvm_host", "-ldpsvm", "-Wl,-rpath," + os.path.join(ROOT, "dpsvm_b200")], check=True)
    return EXE


def _run(exe, *args):
    return subprocess.run([exe] + [str(a) for a in args], capture_output=True, text=True)


def test_populate_data_round_trip(exe, tmp_path):
    x, y = datagen.mnist_like(50, 37, seed=3)
    csv = str(tmp_path / "a.csv")
    datagen.write_csv(csv, x, y)
    r = _run(exe, "parse", csv, 50, 37)
    assert r.returncode == 0
    rows = [l.split(",") for l in r.stdout.splitlines()]
    assert [int(t[0]) for t in rows] == y.tolist()
    assert np.array_equal(np.array([[np.float32(v) for v in t[1:]] for t in rows], np.float32), x)
    r = _run(exe, "parse_rows", csv, 20, 7, 37)                 # shard-only loading
    got = np.array([[np.float32(v) for v in l.split(",")[1:]] for l in r.stdout.splitlines()], np.float32)
    assert np.array_equal(got, x[20:27])


def test_populate_data_edge_cases_follow_the_reference(exe, tmp_path):
    p = tmp_path / "e.csv"
    p.write_text("1,0.5,0.25\n-1,1\n1,2,3,\n")                    # short row keeps zeros; trailing comma adds no cell
    r = _run(exe, "parse", p, 3, 3)
    assert r.returncode == 0
    assert r.stdout.splitlines() == ["1,0.5,0.25,0", "-1,1,0,0", "1,2,3,0"]
    r = _run(exe, "parse", p, 4, 3)                               # file shorter than -x: stoi("") throws in the reference
    assert r.returncode == 10 and "stoi" in r.stdout
    (tmp_path / "bad.csv").write_text("1,abc,2\n")
    r = _run(exe, "parse", tmp_path / "bad.csv", 1, 2)            # stof("abc") throws
    assert r.returncode == 10 and "stof" in r.stdout
    (tmp_path / "long.csv").write_text("1,1,2,3,4\n")
    r = _run(exe, "parse", tmp_path / "long.csv", 1, 2)           # the reference writes past the row; here: an error
    assert r.returncode == 11
    r = _run(exe, "parse", tmp_path / "missing.csv", 1, 2)        # parse.cpp:14-18
    assert r.returncode == 255 and "Couldn't open file" in r.stdout
    (tmp_path / "ws.csv").write_text(" 1, 0.5,1e-3\n")            # stoi / stof skip leading blanks, accept exponents
    r = _run(exe, "parse", tmp_path / "ws.csv", 1, 2)
    assert r.stdout.splitlines() == ["1,0.5,0.00100000005"]


@pytest.mark.skipif(not os.path.exists(SEQ), reason="oracle/_ref/seq not built")
def test_model_writer_is_byte_identical_to_the_reference(exe, tmp_path, oracle, monkeypatch):
    monkeypatch.chdir(tmp_path)                                   # seq.cpp keeps file names in char[60]
    x, y = datagen.mnist_like(300, 24, seed=8)
    datagen.write_csv("t.csv", x, y)
    subprocess.run([SEQ, "-a", "24", "-x", "300", "-f", "t.csv", "-c", "10", "-g", "0.25", "-e", "0.001", "-m", "ref.model"],
                   check=True, capture_output=True)
    res = oracle.train(x, y, 10.0, 0.25, 1e-3)                    # bit-identical to what seq computed (test_oracle_ref)
    res.alpha.tofile("alpha.bin")
    r = _run(exe, "write_model", "t.csv", 300, 24, "alpha.bin", float(res.b), 0.25, 0, "ours.model")
    assert r.returncode == 0
    assert open("ours.model", "rb").read() == open("ref.model", "rb").read()      # seq.cpp:295-321 format
    r = _run(exe, "write_model", "t.csv", 300, 24, "alpha.bin", float(res.b), 0.25, 1, "ours_b.model")
    lines, ref_lines = open("ours_b.model").read().splitlines(), open("ref.model").read().splitlines()
    assert lines[0] == ref_lines[0] and lines[2:] == ref_lines[1:]                 # svmTrainMain.cpp:393-394 adds the b line
    assert abs(float(lines[1]) - float(res.b)) < 1e-5 * max(1.0, abs(float(res.b)))      # default ostream precision: 6 digits
    r = _run(exe, "read_model", "ours_b.model", 24, 1)
    head = r.stdout.splitlines()[0].split()
    assert float(head[0]) == 0.25 and int(head[2]) == int((res.alpha != 0).sum())
    first = r.stdout.splitlines()[1].split(",")
    i0 = int(np.flatnonzero(res.alpha != 0)[0])
    assert int(first[1]) == int(y[i0]) and abs(float(first[0]) - float(res.alpha[i0])) < 1e-5 * max(1.0, float(res.alpha[i0]))
