#!/bin/bash
# build a variant of the library with extra -D flags for K1: tools/k1_variant.sh NAME -DFOO -DBAR  -> build/variants/NAME.so
set -e
name=$1; shift
mkdir -p build/variants
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v "$@" \
  -c edison_b200/csrc/mfcc_q15_kernel.cu -o build/variants/$name.o 2> build/variants/$name.log
grep -A1 "mfcc_q15_kernelILb1" build/variants/$name.log | grep -E "registers|spill" | head -3 || true
grep -B1 -A2 "mfcc_q15_kernelILb1" build/variants/$name.log | grep -E "Used|spill" | head -2
objs=$(ls build/*.o | grep -v mfcc_q15_kernel.o)
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/variants/$name.so build/variants/$name.o $objs \
  -Xlinker --version-script=edison_b200/csrc/exports.map -Xlinker -z,defs -ldl
echo "built build/variants/$name.so"
