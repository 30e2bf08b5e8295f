#!/bin/bash
# on the GPU box: parity tests with B, then alternate A / B benches (tools/ab_build.sh made tmp_ab/lib{A,B}.so)
L=centerfusiondetect3d_b200/libcfd3d_b200.so
cp tmp_ab/libB.so $L
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for i in 1 2 3; do for v in A B; do cp tmp_ab/lib$v.so $L; echo -n "$v cfg2 "; python bench.py --workload cfg2 --steps 1000 --warmup 20 --no-cpu --e2e-steps 2 2>&1 | tail -1 | cut -c128-160; done; done
for w in ${AB_EXTRA:-}; do for v in A B; do cp tmp_ab/lib$v.so $L; echo -n "$v $w "; python bench.py --workload $w --steps 300 --warmup 20 --no-cpu --e2e-steps 2 2>&1 | tail -1 | cut -c126-160; done; done
cp tmp_ab/libB.so $L
