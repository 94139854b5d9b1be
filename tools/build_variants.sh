#!/bin/bash
# Library variants for A/B runs on the GPU box: tools/build_variants.sh name:"-Dflags" ...  -> build/variants/lib_<name>.so
# (same flags as csrc/Makefile plus the variant's defines; run with ISD_LIB_OVERRIDE, tools/gpu_ab_variants.sh)
set -u
cd "$(dirname "$0")/../implisolid_b200/csrc"
OUT=../../build/variants; mkdir -p $OUT
BASE="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC,-O2,-ffp-contract=off -Xptxas -v"
for spec in "$@"; do
  name=${spec%%:*}; flags=${spec#*:}
  ( nvcc $BASE $flags -shared -o $OUT/lib_$name.so geometry_api.cu -lcudart -ldl > $OUT/$name.log 2>&1 && echo "built $name" || echo "FAILED $name" ) &
done
wait
