#!/bin/bash
# build an experimental variant of the library: scripts/build_variant.sh <name> <extra nvcc flags...>
# -> reportree_b200/build/lib_<name>.so (use with RTHC_LIB_PATH=...)
name=$1; shift
cd "$(dirname "$0")/../reportree_b200" || exit 1
F="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -fmad=false"
mkdir -p build/v_$name
for s in api recode collide distance distance_ws mst cut linkage nnchain probe zoom stability; do
  if [ $s = distance ] || [ $s = distance_ws ]; then nvcc $F "$@" -c csrc/$s.cu -o build/v_$name/$s.o & else ln -sf ../$s.o build/v_$name/$s.o; fi
done; wait
nvcc -shared -o build/lib_$name.so build/v_$name/*.o -gencode arch=compute_100a,code=sm_100a -cudart static && echo build/lib_$name.so
