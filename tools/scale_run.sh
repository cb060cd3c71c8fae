#!/bin/bash
# Runs ON a multi-GPU box: bench.py at the given GPU counts for the given workloads; one JSON line each
# into gpurun_out/scale_<workload>_p<N>.json.   usage: scale_run.sh "<N list>" "<workload list>"
set -u
NS="${1:-8}"
WLS="${2:-mnist60k}"
mkdir -p gpurun_out
for wl in $WLS; do
  for n in $NS; do
    out=gpurun_out/scale_${wl}_p${n}.json
    if [ "$n" = "1" ]; then
      timeout 600 python bench.py --workload $wl --steps 8 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out 2> ${out%.json}.err
    else
      timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 \
        bench.py --gpus $n --workload $wl --steps 8 --warmup 3 --no-cpu-baseline --e2e-steps 1 --no-pass-only > $out 2> ${out%.json}.err
    fi
    python - "$out" "$wl" "$n" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("%s P=%s: %.0f iter/s, %.2f us/iter, exchange=%s, e2e=%.0f" % (sys.argv[2], sys.argv[3], d["value"], d["config"]["us_per_iteration"], d["config"]["exchange"], d["e2e"]["value"]))
except Exception as e:
    print("%s P=%s: FAILED %s" % (sys.argv[2], sys.argv[3], e))
PY
  done
done
