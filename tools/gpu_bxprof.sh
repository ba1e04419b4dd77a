#!/bin/bash
# per-kernel time and memory traffic of the benchmark_xcorrs workload: tools/gpu_bxprof.sh <tag> [workload]
tag=$1; wl=${2:-bx}
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,l1tex__t_sector_hit_rate.pct,sm__throughput.avg.pct_of_peak_sustained_elapsed \
   --clock-control none -k regex:big_ -c 40 --csv --log-file gpurun_out/${tag}_bx.csv \
   python bench.py --workload $wl --steps 1 > gpurun_out/${tag}_bx.log 2>&1; echo "ncu rc=$?"
python - <<PY
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/${tag}_bx.csv")) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ii=h.index("ID")
d=collections.OrderedDict()
for r in rows[1:]:
    d.setdefault((r[ii], r[ki][:60]), {})[r[mi]]=r[vi]
for k,v in list(d.items())[-10:]:
    print(k, {a.split('.')[0].replace('__','_'):b for a,b in v.items()})
PY
