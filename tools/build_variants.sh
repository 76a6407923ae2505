#!/bin/bash
# usage: tools/build_variants.sh name1 "-DKF_CFG_F256=8,8,2,4,4" name2 "..." ...  -> variants/lib_<name>.so (parallel)
cd "$(dirname "$0")/.." && mkdir -p variants
while [ $# -gt 1 ]; do
  name=$1; flags=$2; shift 2
  ( nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC -shared $flags \
      hydrogym_b200/csrc/kolmo_b200.cu build/slab.cu.o build/diag.cu.o -o variants/lib_$name.so 2> variants/$name.log || echo "FAILED $name" ) &
done
wait
ls -la variants/*.so
