for d in 0 1 2 3 4 7; do
  (CFD3D_DBG=$d timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:peak_bound -c 12 --csv --log-file gpurun_out/dbg_$d.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-graph > /dev/null 2>&1)
  echo "dbg=$d $(tail -n 6 gpurun_out/dbg_$d.csv | awk -F'","' '{printf "%s ", $NF}')"
done
