#!/bin/bash
# ncu launch list (gpu__time_duration.sum per launch) of a bench.py run: tools/gpu_launches.sh <tag> [bench args]
tag=$1; shift
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
   python bench.py --steps 2 --warmup 3 --no-cpu --no-configs --no-strong "$@" > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${tag}_launches.csv")) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
for r in rows[-16:]: print(r[ki][:70], r[vi])
PY
