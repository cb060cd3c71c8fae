#!/bin/bash
# Runs ON the GPU box (under gpurun): plain bench run, then the ncu launch list and one full capture of
# the dominant kernel for the same command line (default DPSVM_CHUNK: 128 iterations per launch).
# Outputs under gpurun_out/.  $1 = tag, $2 = kernel regex, $3 = extra bench args / env via "env X=Y" prefix in $4.
set -u
TAG="${1:-prof}"
KRE="${2:-smo_resident}"
EXTRA="${3:-}"
ENVP="${4:-}"
mkdir -p gpurun_out
CMD="env $ENVP python bench.py --steps 2 --warmup 3 --e2e-steps 1 --converge 0 --no-cpu-baseline --no-ref-gpu --no-pass-only $EXTRA"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --cache-control none --import-source on -k regex:${KRE} -s 4 -c 1 -o gpurun_out/${TAG}_full -f $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out | grep ${TAG}
