"""What the compiled library is made of, checked without a GPU: the resident loop's kernels must contain the sm_100a
instructions the design rests on (profiles/r02_sass_opcodes.md), the headline kernel must not
spill, and the product must not link a BLAS / Thrust path.  `cuobjdump` comes with the CUDA toolkit the build needs."""
import collections
import os
import re
import shutil
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "..", "dpsvm_b200", "libdpsvm.so")

pytestmark = pytest.mark.skipif(shutil.which("cuobjdump") is None or not os.path.exists(LIB),
                                reason="needs cuobjdump and the built library")


@pytest.fixture(scope="module")
def kernels():
    """{demangled kernel name: Counter of opcodes (predicates stripped)} for every kernel of the library."""
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    assert "sm_100a" in out
    fns, cur = {}, None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = fns.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(.*?);", line)
        if m and cur is not None:
            t = m.group(1).split()
            if t[0].startswith("@"):
                t = t[1:]
            cur[t[0]] += 1
    names = list(fns)
    dem = subprocess.run(["c++filt"] + names, capture_output=True, text=True, check=True).stdout.splitlines()
    return {d.split("(")[0]: fns[n] for d, n in zip(dem, names)}


def _has(ops, pattern):
    return any(re.search(pattern, k) for k in ops)


def test_resident_kernels_stage_the_pair_by_tma_and_reduce_with_redux(kernels):
    res = {k: v for k, v in kernels.items() if "smo_resident_kernel<" in k}
    assert len(res) >= 20                                  # dense geometries + sparse lane counts x column formats
    for name, ops in res.items():
        assert _has(ops, r"^UBLKCP\.S\.G"), name           # cp.async.bulk shared <- global (1-D TMA)
        assert _has(ops, r"^SYNCS\.ARRIVE\.TRANS64") and _has(ops, r"^SYNCS\.PHASECHK\.TRANS64"), name   # its mbarrier
        assert _has(ops, r"^CREDUX\.MIN"), name            # redux.sync candidate minima
        assert _has(ops, r"^STG\.E\.128\.STRONG\.GPU") and _has(ops, r"^LDG\.E\.128\.STRONG\.GPU"), name  # flagged-word slots
        assert _has(ops, r"^STG\.E\.64\.STRONG\.SYS"), name                                              # rank exchange


def test_sparse_sweep_loads_carry_the_l2_priority_in_the_opcode(kernels):
    sparse = {k: v for k, v in kernels.items() if re.search(r"smo_resident_kernel<32, 8, (4|8|16|32), ", k)}
    assert len(sparse) >= 8
    for name, ops in sparse.items():
        assert _has(ops, r"^LDG\.E\.NA\.ELL2\.256"), name
        one_policy = name.rstrip(">").endswith("true")    # <..., C8, KEEP_ALL>
        assert _has(ops, r"^LDG\.E\.NA\.EFL2\.256") != one_policy, name
    for name in ("smo_resident_kernel<32, 8, 8, true, false>", "smo_resident_kernel<32, 8, 8, true, true>"):
        ops = next(v for k, v in sparse.items() if name in k)
        assert ops["PRMT"] >= 8 * 4                        # one permute per 8-bit column distance, four ring slots
        assert not _has(ops, r"^(STL|LDL)"), "the headline kernel spills: %s" % name


def test_predict_is_an_ffma_contraction(kernels):
    ops = next(v for k, v in kernels.items() if "predict_tile_kernel" in k)
    assert ops["FFMA"] >= 1024 and ops["MUFU.EX2"] >= 64   # 8 x 8 per thread x 16-deep tile, expf epilogue


def test_product_links_no_blas_or_thrust():
    ldd = subprocess.run(["ldd", LIB], capture_output=True, text=True).stdout
    assert "cublas" not in ldd.lower() and "nccl" not in ldd.lower()      # NCCL is dlopen'ed on demand (comm.cu)
    syms = subprocess.run("nm -C %s" % LIB, shell=True, capture_output=True, text=True).stdout
    assert "thrust::" not in syms and "cub::" not in syms and "cublas" not in syms.lower()
