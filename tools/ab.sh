#!/bin/bash
# A/B timing of library variants: tools/ab.sh "<lib1> <lib2> ..." "<env settings ; separated>"
# each lib is copied over pydsproutines_b200/libxcorr_b200.so on the GPU box (scratch copy) and bench.py is run.
for lib in $1; do
  cp $lib pydsproutines_b200/libxcorr_b200.so
  IFS=';' read -ra ENVS <<< "${2:-X=1}"
  for e in "${ENVS[@]}"; do
    r=$(env $e python bench.py --steps 5 --warmup 3 --no-cpu 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms'])" 2>&1 | tail -1)
    echo "$lib [$e] ms_per_step kernel_ms: $r" | tee -a gpurun_out/ab.txt
  done
done
