#!/bin/bash
# ncu --set full of one kernel launch of any python command: tools/gpu_ncu_py.sh <tag> <kernel regex> <skip> <python args...>
tag=$1; kre=$2; skip=$3; shift; shift; shift
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$kre -s $skip -c 1 -f -o gpurun_out/${tag} \
   python "$@" > gpurun_out/${tag}_ncu.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/${tag}_ncu.log | cut -c1-300
