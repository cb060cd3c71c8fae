#!/bin/bash
# Runs ON the GPU box (under gpurun): plain bench run, then the ncu launch list and one full capture of
# the dominant kernel for the same command line.  Outputs under gpurun_out/.  $1 = tag, $2 = kernel regex.
set -u
TAG="${1:-prof}"
KRE="${2:-smo_resident}"
mkdir -p gpurun_out
CMD="env DPSVM_CHUNK=16 python bench.py --steps 2 --warmup 3 --iters-per-step 32 --e2e-steps 1 --no-cpu-baseline --no-pass-only"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --cache-control none --import-source on -k regex:${KRE} -s 4 -c 2 -o gpurun_out/${TAG}_full -f $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out | grep ${TAG}
