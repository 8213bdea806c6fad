#!/bin/bash
# GPU call C: parity suite, then phase times of the library and of experimental builds (build_alt/libsvtme_*.so)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/c_tests.log
tail -3 gpurun_out/c_tests.log
for wl in 1080p_p8 4k_p4; do
  timeout 300 python tools/prof_phase.py $wl 3 2>&1 | grep -v "^$" | tail -3
  for v in "$@"; do
    PROF_LIB=build_alt/libsvtme_$v.so timeout 300 python tools/prof_phase.py $wl 3 2>&1 | tail -3
  done
done > gpurun_out/c_phase.log 2>&1
cat gpurun_out/c_phase.log
