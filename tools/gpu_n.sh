timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pingpong or pow2 or golden or warp_layout" 2>&1 | tail -4
for w in ${WL:-n1024 n2048 n4096 n16384}; do python bench.py --workload $w --steps 5 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$w', d['ms_per_step'], d['roofline']['frac'])"; done
