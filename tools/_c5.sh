python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for w in cfg5 cfg2; do timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu --no-pipelined 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w', round(d['ms_per_step']*1000,2),'us', d['parity_check'])"; done
timeout 200 python tools/prof_phases.py cfg5 8 2>&1 | head -16
