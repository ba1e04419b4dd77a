#!/bin/bash
# tools/build_variant.sh <name> <extra nvcc flags...>  -> build/lib_<name>.so (only fastxcorr.cu is rebuilt with the flags)
name=$1; shift
NV="/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v"
mkdir -p build/var_$name
$NV "$@" -c pydsproutines_b200/csrc/fastxcorr.cu -o build/var_$name/fastxcorr.o 2> build/var_$name/ptxas.log || { tail -5 build/var_$name/ptxas.log; exit 1; }
objs=$(ls build/*.o | grep -v fastxcorr.o)
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/lib_$name.so build/var_$name/fastxcorr.o $objs -cudart static
echo -n "$name: "; grep -A2 "fused_fft_kernelINS_7FftPlanILi13ELi256ELi8ELi32ELi32EEELi2ENS_12CutoutLoaderIS2_Lb1EEELb0" build/var_$name/ptxas.log | grep -E "registers" | head -1
