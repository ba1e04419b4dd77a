#!/bin/bash
# ncu --set full of one kernel of a bench.py --workload run: tools/gpu_ncu_cfg.sh <tag> <workload> <kernel regex> [lib]
tag=$1; wl=$2; kre=$3; lib=$4
mkdir -p gpurun_out
[ -n "$lib" ] && cp $lib pydsproutines_b200/libxcorr_b200.so
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$kre -s ${SKIP:-2} -c 1 -f -o gpurun_out/${tag} \
   python bench.py --workload $wl --steps 2 > gpurun_out/${tag}_ncu.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/${tag}_ncu.log | cut -c1-300
