#!/bin/bash
# Variant of the library that differs in how ONE source file is compiled:
#   tools/build_variant.sh <name> <file.cu> -DFLAG=... ->  build/liboz_<name>.so   (other objects are reused from build/csrc)
name=$1; src=$2; shift 2
cd "$(dirname "$0")/../ozone_b200/csrc"
mkdir -p ../../build/var
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
     --expt-relaxed-constexpr "$@" -c $src -o ../../build/var/${src%.cu}_$name.o || exit 1
objs=$(ls ../../build/csrc/*.o | grep -v "/${src%.cu}.o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../build/liboz_$name.so $objs ../../build/var/${src%.cu}_$name.o -lcudart_static -lrt -lpthread -ldl
