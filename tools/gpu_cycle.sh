#!/bin/bash
# one GPU cycle: parity tests, smoke, bench of every workload, phase clocks
TAG=${1:-r2}
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log)
(timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_smoke.log)
for w in cfg2 cfg3 cfg4 cfg5; do (CFD3D_VERBOSE=1 timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu 2> gpurun_out/${TAG}_bench_$w.err > gpurun_out/${TAG}_bench_$w.json; echo "rc=$?" >> gpurun_out/${TAG}_bench_$w.err); done
(timeout 200 python tools/prof_phases.py cfg2 32 4 > gpurun_out/${TAG}_phases_cfg2.txt 2>&1)
tail -n 3 gpurun_out/${TAG}_pytest.log; tail -n 2 gpurun_out/${TAG}_smoke.log
for w in cfg2 cfg3 cfg4 cfg5; do grep "stage plan" gpurun_out/${TAG}_bench_$w.err | sort | uniq -c | tail -2; tail -n 1 gpurun_out/${TAG}_bench_$w.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_bench_$w.json").read().strip().splitlines()[-1])
    print("$w ms/step %.4f value %.0f frac %.4f stage_frac %.4f e2e %.0f parity %s"%(d["ms_per_step"],d["value"],d["roofline"]["frac"],d["roofline"]["stage"]["frac"],d["e2e"]["value"],d["parity_check"]))
except Exception as e: print("$w: no line", e)
PY
done
tail -n 40 gpurun_out/${TAG}_phases_cfg2.txt | head -60
