#!/bin/bash
# A/B on one box: tmp_ab/libA.so = the committed tree (HEAD), tmp_ab/libB.so = the working tree.  Then on the GPU:
#   for v in A B; do cp tmp_ab/lib$v.so centerfusiondetect3d_b200/libcfd3d_b200.so; python bench.py ...; done
set -e
cd "$(dirname "$0")/.."
python -c "from centerfusiondetect3d_b200 import build; build.build()"
mkdir -p tmp_ab
cp centerfusiondetect3d_b200/libcfd3d_b200.so tmp_ab/libB.so
git stash -q
python -c "from centerfusiondetect3d_b200 import build; build.build()" || { git stash pop -q; exit 1; }
cp centerfusiondetect3d_b200/libcfd3d_b200.so tmp_ab/libA.so
git stash pop -q
cp tmp_ab/libB.so centerfusiondetect3d_b200/libcfd3d_b200.so
echo "A = HEAD, B = working tree"
