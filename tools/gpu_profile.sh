#!/bin/bash
# Runs ON the GPU box (under gpurun): plain bench run, then the ncu launch list and one full
# capture of the pass kernel for the same command line.  Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
CMD="env DPSVM_CHUNK=32 python bench.py --steps 2 --warmup 3 --iters-per-step 32 --e2e-steps 1 --no-cpu-baseline --no-pass-only"
$CMD > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || { echo "plain run failed"; tail -5 gpurun_out/prof_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:smo_pass -s 40 -c 3 -o gpurun_out/pass_full -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out
