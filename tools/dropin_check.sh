#!/bin/bash
# Drop-in proof: build the UNMODIFIED reference driver (/root/reference/svmTrainMain.cpp) against
# THIS repo's include/svmTrain.h + libdpsvm_host.so, with a single-rank mpi.h stand-in.
# The reference source is compiled from a scratch copy outside the repo (a quoted #include would
# otherwise pick the reference's own svmTrain.h from its directory); nothing is copied into the repo.
# Output: build/svmTrain_refmain (travels to the GPU box with gpurun).
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
REF="${REF:-/root/reference}"
[ -f "$REF/svmTrainMain.cpp" ] || { echo "no reference checkout at $REF"; exit 3; }
TMP="$(mktemp -d)"
trap 'rm -rf "$TMP"' EXIT
cp "$REF/svmTrainMain.cpp" "$TMP/"
mkdir -p "$ROOT/build"
g++ -O2 -std=c++11 -w -I"$ROOT/include" -I"$ROOT/tools/dropin" -o "$ROOT/build/svmTrain_refmain" "$TMP/svmTrainMain.cpp" \
    -L"$ROOT/dpsvm_b200" -ldpsvm_host -ldpsvm -Wl,-rpath,"$ROOT/dpsvm_b200" -Wl,-rpath,'$ORIGIN/../dpsvm_b200'
echo "built $ROOT/build/svmTrain_refmain from unmodified $REF/svmTrainMain.cpp"
