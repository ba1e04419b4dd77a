for lib in build/lib_noblue.so build/cur.so; do cp $lib pydsproutines_b200/libxcorr_b200.so; echo "== $lib"; python tools/bench_paths.py blue1000 blue2000 2>&1 | tail -2 | cut -c1-140; done
