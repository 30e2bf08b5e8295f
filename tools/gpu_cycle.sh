(timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1_pytest.log)
for w in $WLS; do (timeout 400 python bench.py --workload $w --steps 300 --warmup 20 --no-cpu > gpurun_out/r1_bench_$w.log 2>&1; echo "rc=$?" >> gpurun_out/r1_bench_$w.log); done
(timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r1_launches_cfg2.csv python bench.py --steps 5 --warmup 3 --no-cpu --no-graph > gpurun_out/r1_ncu_cfg2.log 2>&1)
tail -n 3 gpurun_out/r1_pytest.log
