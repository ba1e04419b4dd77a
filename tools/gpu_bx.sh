for mb in 96 48 200 1024 4096; do echo "== XC_BIG_BATCH_MB=$mb"; XC_BIG_BATCH_MB=$mb python bench.py --workload bx --steps 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'], d['vs_baseline'])"; done
python bench.py --workload bx10 --steps 3 2>&1 | tail -1 | cut -c1-1200
