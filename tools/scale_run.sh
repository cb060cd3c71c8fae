#!/bin/bash
# Runs ON a multi-GPU box: bench.py at the given GPU counts for the given workloads; one JSON line each
# into gpurun_out/<tag>_scale_<workload>_p<N>.json.   usage: scale_run.sh "<N list>" "<workload list>" [tag] [extra bench args]
set -u
NS="${1:-8}"
WLS="${2:-mnist60k}"
TAG="${3:-r02}"
EXTRA="${4:-}"
mkdir -p gpurun_out
for wl in $WLS; do
  for n in $NS; do
    out=gpurun_out/${TAG}_scale_${wl}_p${n}.json
    if [ "$n" = "1" ]; then
      timeout 900 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu $EXTRA > $out 2> ${out%.json}.err
    else
      timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 \
        bench.py --gpus $n --workload $wl --steps 10 --warmup 3 --no-cpu-baseline $EXTRA > $out 2> ${out%.json}.err
    fi
    python - "$out" "$wl" "$n" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    par = d.get("parity") or {}
    ttc = d.get("time_to_converge") or {}
    big = d.get("time_to_converge_hbm_cache") or {}
    print("%s P=%s: %.0f iter/s, %.2f us/iter, exchange=%s, e2e=%.0f, bit_identical=%s, n1=%s, speedup_vs_n1=%s, converge=%.3fs (%s it), hbm-cache converge=%s" % (
        sys.argv[2], sys.argv[3], d["value"], d["run"]["us_per_iteration"], d["run"]["exchange"], d["e2e"]["value"], par.get("bit_identical"),
        round(d["n1"]["value"]) if d.get("n1") else None, round(d["speedup_vs_n1"], 3) if d.get("speedup_vs_n1") else None,
        ttc.get("seconds", float("nan")), ttc.get("iterations"), big.get("seconds")))
except Exception as e:
    print("%s P=%s: FAILED %s" % (sys.argv[2], sys.argv[3], e))
PY
  done
done
