#!/bin/bash
# on the GPU box: parity tests, then the same library benched with and without an environment switch, alternating
#   tools/ab_env.sh CFD3D_NO_TLBWARM "cfg2 cfg3"
VAR=$1; WLS=${2:-cfg2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for w in $WLS; do for i in 1 2 3; do
  echo -n "$w default   "; python bench.py --workload $w --steps 1000 --warmup 20 --no-cpu --e2e-steps 2 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%.2f us' % (d['ms_per_step']*1e3))"
  echo -n "$w $VAR=1 "; env $VAR=1 python bench.py --workload $w --steps 1000 --warmup 20 --no-cpu --e2e-steps 2 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%.2f us' % (d['ms_per_step']*1e3))"
done; done
