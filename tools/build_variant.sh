#!/bin/bash
# tools/build_variant.sh NAME [-Dmacro ...]: tuning build of libduodec with macro overrides -> duodecoding_b200/build/variants/NAME.so
set -e
name=$1; shift
mkdir -p duodecoding_b200/build/variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC "$@" -c -o /tmp/sl_$name.o duodecoding_b200/csrc/stream_launch.cu
nvcc -gencode arch=compute_100a,code=sm_100a --shared -o duodecoding_b200/build/variants/$name.so duodecoding_b200/build/c_abi.o /tmp/sl_$name.o
echo built $name
