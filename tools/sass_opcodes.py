#!/usr/bin/env python
"""Opcode evidence from the built objects: `python tools/sass_opcodes.py > profiles/r02_sass_opcodes.md`.

Runs `cuobjdump -sass` on build/smo_resident.o and build/predict.o (what `make lib` leaves), prints the signature
instructions (TMA bulk copy, mbarrier, redux, the stream loads with their L2 priority, the flagged-word stores and
polls, the named barrier) with counts, then the whole opcode histogram of the kernels the bench line runs.
"""
import collections
import re
import subprocess
import sys

KERNELS = [
    ("build/smo_resident.o", "smo_resident_kernel<32, 8, 8, true, false>", "sparse stream, 8 lanes per row, 8-bit column distances, two eviction policies (a stream about the size of L2: BASELINE configs[1])"),
    ("build/smo_resident.o", "smo_resident_kernel<32, 8, 8, true, true>", "the same with one policy (a stream well inside L2: the shards of N > 1)"),
    ("build/smo_resident.o", "smo_resident_kernel<32, 7, 0, false, false>", "dense rows of 784 floats"),
    ("build/predict.o", "predict_tile_kernel", "fp32 FFMA contraction: 128 x 128 tile, 8 x 8 per thread"),
]
SIGNATURE = re.compile(r"UBLKCP|SYNCS|CREDUX|LDG\.E\.NA|\.STRONG\.|BAR\.(ARV|SYNC\.DEFER_BLOCKING 0x1)|PRMT|UTMA|REDUX")


def functions(obj):
    names = subprocess.run("cuobjdump -sass %s | grep 'Function :' | awk '{print $3}'" % obj, shell=True, capture_output=True, text=True).stdout.split()
    dem = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    return dict(zip(dem, names))


def sass(obj, mangled):
    out = subprocess.run(["cuobjdump", "-sass", "-fun", mangled, obj], capture_output=True, text=True).stdout
    ins = []
    for line in out.splitlines():
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(.*?);", line)
        if m:
            ins.append(m.group(1).strip())
    return ins


def opcode(i, with_pred):
    t = i.split()
    pred = ""
    if t[0].startswith("@"):
        pred, t = t[0] + " ", t[1:]
    return (pred if with_pred else "") + t[0]


def main():
    head = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    print("# SASS opcode evidence (round 2, kernels as of commit %s)\n" % head)
    print("`python tools/sass_opcodes.py`: `cuobjdump -sass build/smo_resident.o` / `build/predict.o` (nvcc 12.9, `-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo`).")
    print("Signature instructions first (count, instruction), then the full opcode histogram of each kernel.\n")
    print("""| what | SASS |
|---|---|
| 1-D TMA bulk copy of the selected pair's two rows into shared memory | `UBLKCP.S.G` + `SYNCS.ARRIVE.TRANS64` / `SYNCS.PHASECHK.TRANS64.TRYWAIT` (mbarrier) |
| 256-bit stream loads, L2 eviction priority in the opcode, L1 no-allocate | `LDG.E.NA.ELL2.256` (evict_last), `.EFL2.256` (evict_first) |
| 8-bit column distances: one 64-bit load per lane and chunk, one permute + one multiply-add per entry | `LDG.E.NA.64.CONSTANT`, `PRMT ... 0x444k`, `IMAD` |
| candidate reductions, cache victim | `CREDUX.MIN` (`redux.sync`) |
| flagged-word slots that replace the grid barrier | `STG.E.128.STRONG.GPU` (two words per store), `LDG.E.128.STRONG.GPU` (polling) |
| flagged-word candidate exchange between GPUs (peer memory) | `STG.E.64.STRONG.SYS`, `LDG.E.64.STRONG.SYS` |
| eta handed over to the sweeping warps through named barrier 1 | `BAR.ARV 0x1, 0x200`, `BAR.SYNC.DEFER_BLOCKING 0x1, 0x200` |
""")
    cache = {}
    for obj, want, what in KERNELS:
        fn = cache.setdefault(obj, functions(obj))
        hits = [d for d in fn if want in d]
        if not hits:
            print("## %s: not in %s\n" % (want, obj))
            continue
        dem = hits[0]
        ins = sass(obj, fn[dem])
        print("## `%s`\n\n%s; %d instructions.\n" % (dem.split("(")[0], what, len(ins)))
        sig = collections.Counter(opcode(i, True) + (" " + " ".join(i.split()[-2:]) if "BAR." in i else "") for i in ins if SIGNATURE.search(i))
        if sig:
            print("Signature instructions:\n\n```")
            for k, v in sig.most_common():
                print("%7d %s" % (v, k))
            print("```\n")
        print("All opcodes:\n\n```")
        for k, v in collections.Counter(opcode(i, False) for i in ins).most_common():
            print("%7d %s" % (v, k))
        print("```\n")


if __name__ == "__main__":
    sys.exit(main())
