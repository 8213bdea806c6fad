#!/bin/bash
# bench_variants.sh NAME... : phase times of each experimental library of build_alt/ (tools/prof_step.py, GPU box)
cd "$(dirname "$0")/.."
L=svt-av1-mod-by-patman_b200/csrc/libsvtme_b200.so
cp $L /tmp/libsvtme_keep.so
for v in "$@"; do
  cp build_alt/libsvtme_$v.so $L
  python tools/prof_step.py ${BENCH_ARGS} > gpurun_out/var_$v.txt 2>&1; echo $v; cat gpurun_out/var_$v.txt
done
cp /tmp/libsvtme_keep.so $L
