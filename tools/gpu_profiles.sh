#!/bin/bash
# one GPU call: everything profiles/<round>_* is made of (tools/make_profiles.py turns it into the tracked files)
TAG=${1:-r2}
mkdir -p gpurun_out
for w in cfg2 cfg3 cfg4 cfg5; do (timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu 2> gpurun_out/${TAG}_bench_$w.err > gpurun_out/${TAG}_bench_$w.json); done
(timeout 300 python bench.py > gpurun_out/${TAG}_bench_default.json 2> gpurun_out/${TAG}_bench_default.err)
(timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_driver.json 2>/dev/null)
(timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_reference.json 2>/dev/null)
(timeout 200 python tools/prof_phases.py cfg2 32 4 > gpurun_out/${TAG}_phases.txt 2>&1)
(COLD=1 timeout 200 python tools/prof_phases.py cfg2 32 4 > gpurun_out/${TAG}_phases_cold.txt 2>&1)
(timeout 200 ./tools/gather_floor > gpurun_out/${TAG}_gather_floor.txt 2>&1)
(timeout 600 bash tools/plan_sweep.sh "cfg2 cfg3 cfg4 cfg5" "1 2 4 8 16" > gpurun_out/${TAG}_plan_sweep.txt 2>&1)
# ncu: launch lists (durations only), then one full capture per workload -- numbers under a profiler are never bench values
for w in cfg2 cfg4; do
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_launches_$w.csv python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --no-graph --e2e-steps 2 > /dev/null 2>&1
done
for w in cfg2 cfg4 cfg5; do
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:fusion_stage -s 6 -c 1 -o gpurun_out/${TAG}_stage_full_$w -f python bench.py --workload $w --steps 5 --warmup 3 --no-cpu --no-graph --e2e-steps 2 > gpurun_out/${TAG}_ncu_$w.log 2>&1
done
ls -la gpurun_out | tail -30
