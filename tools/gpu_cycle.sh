#!/bin/bash
# one GPU cycle: parity tests, smoke, bench of every workload
TAG=${1:-r1c}
mkdir -p gpurun_out
(timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log)
(timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_smoke.log)
for w in cfg2 cfg3 cfg4 cfg5; do (CFD3D_VERBOSE=1 timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu 2> gpurun_out/${TAG}_bench_$w.err > gpurun_out/${TAG}_bench_$w.json; echo "rc=$?" >> gpurun_out/${TAG}_bench_$w.err); done
tail -n 3 gpurun_out/${TAG}_pytest.log; tail -n 2 gpurun_out/${TAG}_smoke.log
for w in cfg2 cfg3 cfg4 cfg5; do grep "stage plan" gpurun_out/${TAG}_bench_$w.err | sort | uniq -c | tail -2; tail -n 1 gpurun_out/${TAG}_bench_$w.err; cut -c1-220 gpurun_out/${TAG}_bench_$w.json; done
