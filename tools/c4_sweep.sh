#!/bin/bash
# C4 option sweep (developer tool): prints value / refactor / solve per option set
for opts in "" "--option max_super=128" "--option max_super=512" "--option blocked_lu_min_dim=64" "--option blocked_lu_min_dim=160" "--option max_super=128 --option blocked_lu_min_dim=64" "--option max_super=512 --option blocked_lu_min_dim=160" "--option leaf_size=8" "--option unit_threads=128"; do
  python bench.py --workload grid2d --steps 4 --warmup 3 --no-cpu-baseline --no-e2e $opts 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$opts', round(d['value']), 'refactor', round(d['phases']['refactor']['us']), 'solve', round(d['phases']['solve']['us']), 'launches/iter', d['gpu_launches']/4, 'levels', d['factor']['n_levels'], 'nnz', d['factor']['nnz_factor'])"
done
