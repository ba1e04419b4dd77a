#!/bin/bash
# tools/build_pp.sh <name> [extra nvcc flags]: rebuilds only search_pp.cu with the flags and links build/lib_<name>.so
# from the other objects of the last `make` (A/B variants of the ping-pong search kernel).
name=$1; shift
NV="/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v -DXC_EXPERIMENTS"
mkdir -p build/var_$name
$NV "$@" -c pydsproutines_b200/csrc/search_pp.cu -o build/var_$name/search_pp.o 2> build/var_$name/ptxas.log || { tail -5 build/var_$name/ptxas.log; exit 1; }
objs=$(ls build/*.o | grep -v search_pp.o)
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/lib_$name.so build/var_$name/search_pp.o $objs -cudart static
echo -n "$name: "; grep "Used" build/var_$name/ptxas.log | tr '\n' ' '; grep -c "spill stores" build/var_$name/ptxas.log; grep "bytes spill" build/var_$name/ptxas.log | head -2
