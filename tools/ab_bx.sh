#!/bin/bash
# A/B of library variants on the benchmark_xcorrs workloads: tools/ab_bx.sh <tag> <lib...>
tag=$1; shift
mkdir -p gpurun_out
for lib in "$@"; do
  cp $lib pydsproutines_b200/libxcorr_b200.so
  for w in ${WL:-bx bx10}; do
    echo -n "$lib $w " | tee -a gpurun_out/${tag}_abbx.txt
    python bench.py --workload $w --steps 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'])" | tee -a gpurun_out/${tag}_abbx.txt
  done
done
