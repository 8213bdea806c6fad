#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dg_detector.py tests/test_stats.py -m gpu -x -q > gpurun_out/f_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/f_tests.log
tail -15 gpurun_out/f_tests.log
