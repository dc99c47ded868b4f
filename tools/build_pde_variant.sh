#!/bin/bash
# Variant of the library that differs only in how oz_pde.cu is compiled:
#   tools/build_pde_variant.sh <name> -DOZ_PDE_CPT=3 ...   ->  build/liboz_<name>.so
name=$1; shift
cd "$(dirname "$0")/../ozone_b200/csrc"
mkdir -p ../../build/var
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
     --expt-relaxed-constexpr "$@" -c oz_pde.cu -o ../../build/var/oz_pde_$name.o -Xptxas -v 2>&1 | grep -A2 "pde_step_kernelIdLi4\|pde_adjoint_step_kernelIdLi4" | grep Used
objs=$(ls ../../build/csrc/*.o | grep -v oz_pde.o)
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../build/liboz_$name.so $objs ../../build/var/oz_pde_$name.o -lcudart_static -lrt -lpthread -ldl
