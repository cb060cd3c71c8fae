#!/bin/bash
# Runs ON the GPU box: A/B two builds of libdpsvm.so (DPSVM_LIB) on the default bench, interleaved.
# usage: ab_bench.sh <libA> <libB> [repeats]
A="$1"; B="$2"; N="${3:-3}"
for i in $(seq 1 $N); do
  for L in "$A" "$B"; do
    DPSVM_LIB=$L timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 1 --no-pass-only 2>/dev/null | \
      python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$L', round(d['value']), round(d['config']['us_per_iteration'],2), round(d['e2e']['value']))"
  done
done
