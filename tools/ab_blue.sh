for lib in "$@"; do cp $lib pydsproutines_b200/libxcorr_b200.so; echo "== $lib"; python tools/bench_paths.py blue3000 2>&1 | tail -1 | cut -c1-140; done
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pingpong_bluestein" 2>&1 | tail -2
