#!/bin/bash
# one GPU cycle: parity tests, bench (cfg2 default), launch list + full ncu capture of the stage kernel
TAG=${1:-r1c}
mkdir -p gpurun_out
(timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log)
(CFD3D_VERBOSE=1 timeout 300 python bench.py --steps 2000 --warmup 50 > gpurun_out/${TAG}_bench_cfg2.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_bench_cfg2.log)
tail -n 3 gpurun_out/${TAG}_pytest.log
tail -n 2 gpurun_out/${TAG}_bench_cfg2.log | cut -c1-2500
