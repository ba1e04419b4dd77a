#!/bin/bash
# launch list (all kernels) of the last step of the benchmark_xcorrs workload
tag=$1; wl=${2:-bx}
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_bxall.csv \
   python bench.py --workload $wl --steps 1 > gpurun_out/${tag}_bxall.log 2>&1; echo "ncu rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${tag}_bxall.csv")) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
for r in rows[-26:]: print(r[ki][:60], r[vi])
PY
