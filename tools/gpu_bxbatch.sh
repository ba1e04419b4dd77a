cp build/lib_bigexp.so pydsproutines_b200/libxcorr_b200.so
for mb in 64 96 128 256 512 1024 2048; do echo -n "XC_BIG_BATCH_MB=$mb "; XC_BIG_BATCH_MB=$mb python bench.py --workload bx --steps 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'])"; done
for mb in 512 1024 2048 4096; do echo -n "bx10 XC_BIG_BATCH_MB=$mb "; XC_BIG_BATCH_MB=$mb python bench.py --workload bx10 --steps 3 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'])"; done
