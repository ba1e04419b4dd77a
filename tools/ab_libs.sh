#!/bin/bash
# A/B of library variants on the GPU box: tools/ab_libs.sh <tag> <lib1> <lib2> ...   (each lib replaces the in-tree .so of
# the scratch copy; layouts b2 and pp are timed and compared by tools/staged_check.py)
tag=$1; shift
mkdir -p gpurun_out
for lib in "$@"; do
  cp $lib pydsproutines_b200/libxcorr_b200.so
  echo "== $lib" | tee -a gpurun_out/${tag}_ab.txt
  timeout 600 python tools/staged_check.py b2 pp 2>&1 | grep '"pp"' | tee -a gpurun_out/${tag}_ab.txt
done
