#!/bin/bash
# A/B of the GroupXcorrFFT kernels: round-1 ELL kernel vs ping-pong with B = 2, 3, 4 thread groups per SM
mkdir -p gpurun_out
for b in 3; do
  cp build/lib_gp$b.so pydsproutines_b200/libxcorr_b200.so
  timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -x -q -k "group or c3 or C3" 2>&1 | tail -3
done
echo "== ell"; XC_GROUP_LAYOUT=ell python bench.py --workload c3 --steps 5 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'])"
for b in 2 3 4; do
  cp build/lib_gp$b.so pydsproutines_b200/libxcorr_b200.so
  echo "== pp B=$b"; python bench.py --workload c3 --steps 5 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'])"
done
