#!/bin/bash
# Kernel-tuning experiments: an alternative build of some source files linked with the regular objects.
# usage: build_variant.sh <name> <source.cu>... -- <-D flags...>   -> omldm_b200/lib/libvar_<name>.so
# (load it with OMLDM_KERNELS_LIB=omldm_b200/lib/libvar_<name>.so; never committed, never the default)
set -e
name=$1; shift
srcs=()
while [ $# -gt 0 ] && [ "$1" != "--" ]; do srcs+=("$1"); shift; done
[ "$1" == "--" ] && shift
python -m omldm_b200.ops.build > /dev/null
mkdir -p build/var
objs=$(ls build/obj/*.o)
extra=()
for src in "${srcs[@]}"; do
  stem=$(basename $src .cu)
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr \
       -I include -I csrc "$@" -c $src -o build/var/${stem}_$name.o
  objs=$(echo "$objs" | grep -v "/$stem.o")
  extra+=("build/var/${stem}_$name.o")
done
nvcc -shared -cudart shared -o omldm_b200/lib/libvar_$name.so $objs "${extra[@]}" 2>/dev/null
echo omldm_b200/lib/libvar_$name.so
