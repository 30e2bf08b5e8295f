# cumulative in-pipeline step time after n kernels (graph replay, rotating inputs)
for n in 0 1 2 3 4 5; do
  echo "stop_after=$n $(CFD3D_STOP_AFTER=$n timeout 200 python bench.py --workload ${WL:-cfg2} --steps 400 --warmup 20 --no-cpu --e2e-steps 3 2>&1 | grep -o '"ms_per_step": [0-9.e-]*' | head -1)"
done
