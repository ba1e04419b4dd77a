#!/bin/bash
# A/B of the GroupXcorrFFT kernels: round-1 ELL kernel vs ping-pong with B thread groups per SM
# usage: tools/gpu_gp.sh "<B values>"
mkdir -p gpurun_out
first=1
for b in $1; do
  cp build/lib_gp$b.so pydsproutines_b200/libxcorr_b200.so
  if [ $first = 1 ]; then timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -x -q -k "group or c3 or C3" 2>&1 | tail -3; first=0; fi
  echo "== pp B=$b"; python bench.py --workload c3 --steps 5 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'])"
done
