#!/bin/bash
# GPU call E: statistics tests + timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_stats.py -m gpu -x -q > gpurun_out/e_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/e_tests.log
tail -15 gpurun_out/e_tests.log
for wl in 1080p_p8 4k_p4; do timeout 300 python tools/prof_stats.py $wl 5 2>&1 | tail -2; done > gpurun_out/e_stats.log 2>&1
cat gpurun_out/e_stats.log
