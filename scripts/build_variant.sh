#!/bin/bash
# Kernel-tuning experiments: an alternative build of ONE source file linked with the regular objects.
# usage: build_variant.sh <name> <source.cu> <-D flags...>   -> omldm_b200/lib/libvar_<name>.so  (load with OMLDM_KERNELS_LIB)
set -e
name=$1; src=$2; shift 2
python -m omldm_b200.ops.build > /dev/null
mkdir -p build/var
stem=$(basename $src .cu)
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr \
     -I include -I csrc "$@" -c $src -o build/var/${stem}_$name.o
objs=$(ls build/obj/*.o | grep -v "/$stem.o")
nvcc -shared -cudart shared -o omldm_b200/lib/libvar_$name.so $objs build/var/${stem}_$name.o
echo omldm_b200/lib/libvar_$name.so
