#!/bin/bash
# A/B of option settings on ONE box: bash profiles/tools/ab.sh <workload> "<opts A>" "<opts B>" ...
# prints value and phase times per variant; each variant is run twice, interleaved.
wl=$1; shift
for rep in 1 2; do
  for v in "$@"; do
    o=""; for kv in $v; do o="$o --option $kv"; done
    timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-e2e $o 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$wl [$v]', round(d['value'],2), {k:round(x['us'],1) for k,x in d['phases'].items()}, 'frac', d['roofline'] and round(d['roofline']['frac'],4))"
  done
done
