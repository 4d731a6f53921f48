#!/bin/bash
# tools/build_variant.sh NAME FILE.cu "-DFOO=1 ..." : links flexflow_b200/libffb200_NAME.so with FILE.cu recompiled
# with the extra flags and every other object of the regular build (run `python -m flexflow_b200.build` first).
set -e
cd "$(dirname "$0")/../flexflow_b200"
name=$1; src=$2; flags=$3
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $flags -c csrc/$src -o _obj/variant_$name.o
objs=$(ls _obj/*.o | grep -v "_obj/variant_" | grep -v "_obj/${src%.cu}.o")
nvcc -shared -o libffb200_$name.so $objs _obj/variant_$name.o -lcudart -ldl
echo built libffb200_$name.so
