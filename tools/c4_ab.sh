#!/bin/bash
# C4 A/B on one box (developer tool): bash tools/c4_ab.sh "<k=v k=v>" "<k=v ...>" ...  -> value / refactor / solve per variant
for v in "$@"; do
  o=""; for kv in $v; do o="$o --option $kv"; done
  timeout 300 python bench.py --workload grid2d --steps 10 --warmup 3 --no-cpu-baseline --no-e2e $o 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('[$v]', round(d['value']), 'refactor', round(d['phases']['refactor']['us']), 'solve', round(d['phases']['solve']['us']), 'launches/iter', d['gpu_launches']/10, 'parity', d.get('parity'))"
done
