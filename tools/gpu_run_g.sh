#!/bin/bash
# GPU call G: phase times of the library and of experimental builds, no tests
mkdir -p gpurun_out
for wl in 1080p_p8 4k_p4; do
  timeout 300 python tools/prof_phase.py $wl 3 2>&1 | grep -v "^$" | tail -3 | head -2
  for v in "$@"; do
    echo "--- $v"; PROF_LIB=build_alt/libsvtme_$v.so timeout 300 python tools/prof_phase.py $wl 3 2>&1 | tail -3 | head -2
  done
done > gpurun_out/g_phase.log 2>&1
cat gpurun_out/g_phase.log
