#!/bin/bash
# ncu --set full capture of one kernel of a bench.py run: tools/gpu_ncu.sh <tag> <kernel regex> [bench args...]
tag=$1; shift; kre=$1; shift
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-configs --no-strong "$@" > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
cat gpurun_out/${tag}_bench.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$kre -s 3 -c 1 -f -o gpurun_out/${tag} \
   python bench.py --steps 2 --warmup 3 --no-cpu --no-configs --no-strong "$@" > gpurun_out/${tag}_ncu.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/${tag}_ncu.log
