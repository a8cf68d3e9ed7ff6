#!/bin/bash
# Builds a kernel-variant copy of the product library for A/B timing on the GPU box:
#   profiles/build_variant.sh <name> [-DFLAG ...]  ->  starkpack_winterfell_b200/variants/libwinter_gpu_<name>.so
# (selected at run time with WG_LIB_PATH, see starkpack_winterfell_b200/_lib.py). Not part of the product build.
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
src=$root/starkpack_winterfell_b200/csrc
out=$root/starkpack_winterfell_b200/variants
tmp=/tmp/wg_variant_$name
mkdir -p $out $tmp
FLAGS="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -I$root/include $*"
for f in ntt_gl64 ntt_f128 api; do
  ( cd $src && nvcc $FLAGS -c $f.cu -o $tmp/$f.o ) &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libwinter_gpu_$name.so $tmp/ntt_gl64.o $tmp/ntt_f128.o $tmp/api.o -lcudart
ls -la $out/libwinter_gpu_$name.so
