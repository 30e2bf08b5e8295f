#!/bin/bash
# on the GPU box: step time of a workload under every cluster size the planner could pick (CFD3D_FORCE_CS)
#   tools/plan_sweep.sh "cfg3 cfg4 cfg5" "1 2 4 8 16"
for w in ${1:-cfg4}; do for cs in ${2:-1 2 4 8 16}; do
  echo -n "$w cs=$cs  "; CFD3D_FORCE_CS=$cs python bench.py --workload $w --steps 200 --warmup 10 --no-cpu --e2e-steps 2 2>/dev/null | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); p=d['run']['plan']; print('%.1f us  clusters %d rounds %d resident %d parity %s' % (d['ms_per_step']*1e3, p['n_clusters'], p['rounds'], p['resident'], d['parity_check']))
except Exception as e: print('failed', e)"
done; done
