#!/bin/bash
# build_variant.sh NAME [-DFLAG ...] : experimental build of the library into build_alt/libsvtme_NAME.so
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p build_alt/$name
C=svt-av1-mod-by-patman_b200/csrc
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-fopenmp -Xptxas -v -Iinclude"
for f in svtme_api.cu svtme_pyramid.cu svtme_stage.cu svtme_testabi.cu svtme_noalloc.cu svtme_stats.cu svtme_params.cpp; do
  ( nvcc $FLAGS "$@" -x cu -c $C/$f -o build_alt/$name/${f%.*}.o 2> build_alt/$name/${f%.*}.log || { cat build_alt/$name/${f%.*}.log; exit 1; } ) &
done
wait
nvcc -shared -o build_alt/libsvtme_$name.so build_alt/$name/*.o -lcudart -lgomp
echo built build_alt/libsvtme_$name.so
