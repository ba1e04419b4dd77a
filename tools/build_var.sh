#!/bin/bash
# tools/build_var.sh <name> <source stem> [extra nvcc flags...] -> build/lib_<name>.so: only csrc/<stem>.cu is rebuilt with
# the flags, the other objects come from the last `make` (A/B variants of one translation unit)
name=$1; stem=$2; shift; shift
NV="/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v"
mkdir -p build/var_$name
$NV "$@" -c pydsproutines_b200/csrc/$stem.cu -o build/var_$name/$stem.o 2> build/var_$name/ptxas.log || { tail -5 build/var_$name/ptxas.log; exit 1; }
objs=$(ls build/*.o | grep -v "build/$stem.o")
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/lib_$name.so build/var_$name/$stem.o $objs -cudart static -ldl
echo "$name: spills $(grep -c 'spill stores' build/var_$name/ptxas.log) nonzero $(grep 'spill stores' build/var_$name/ptxas.log | grep -vc ' 0 bytes spill stores')"
