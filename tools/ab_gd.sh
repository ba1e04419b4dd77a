for lib in "$@"; do cp $lib pydsproutines_b200/libxcorr_b200.so; echo "== $lib"; python tools/bench_paths.py group128 group256 group64 2>&1 | tail -4 | cut -c1-120; done
