"""bench.py's reference arm on CPU (the oracle-side leg; no GPU needed): one JSON line with the
contract's keys, rank > 0 exits without work."""
import json
import os
import subprocess
import sys

from conftest import ROOT


def _run(env_extra=None):
    env = dict(os.environ)
    env.update(env_extra or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--rows", "600",
                           "--steps", "2", "--warmup", "1"], capture_output=True, text=True, env=env, timeout=300)


def test_reference_arm_prints_one_contract_line():
    r = _run()
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "smo_iterations_per_s" and d["unit"] == "iter/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] == 1
    assert d["cpu_baseline"]["value"] == d["value"] and d["e2e"]["value"] == d["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"]


def test_reference_arm_other_ranks_exit_quietly():
    r = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_algorithmic_bytes_are_survey_8d():
    """roofline.achieved is built from these: 4 n d + 20 n per iteration, + 8 n of cache-line stores when
    the kernel-row cache is on; the sparse stream's own basis counts 6 B per non-zero, 5 B with 8-bit column distances."""
    sys.path.insert(0, ROOT)
    import bench
    n, d, nnz = 60000, 784, 18223215
    assert bench.algorithmic_bytes(n, d, 0) == 4 * n * d + 20 * n
    assert bench.algorithmic_bytes(n, d, 10) == 189840000                    # the figure DESIGN.md / BASELINE.md quote
    assert bench.algorithmic_bytes(n, d, 10, nnz) == 6 * nnz + 28 * n
    assert bench.algorithmic_bytes(n, d, 10, nnz, sweep_bytes=int(5 * nnz * 1.125)) == 5 * nnz + 28 * n
    assert bench.algorithmic_bytes(n, d, 10, nnz, sweep_bytes=int(6 * nnz * 1.125)) == 6 * nnz + 28 * n
