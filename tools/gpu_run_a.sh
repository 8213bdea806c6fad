#!/bin/bash
# GPU call A: whole GPU suite, the no-alloc test alone in a fresh process, bench at 1080p and 4K.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/a_smi.txt 2>&1
nproc >> gpurun_out/a_smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/a_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/a_tests.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/a_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/a_noalloc_alone.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/a_bench_1080p.json 2> gpurun_out/a_bench_1080p.err; echo "rc=$?" >> gpurun_out/a_bench_1080p.err
timeout 900 python bench.py --workload 4k_p4 --steps 3 --warmup 3 > gpurun_out/a_bench_4k.json 2> gpurun_out/a_bench_4k.err; echo "rc=$?" >> gpurun_out/a_bench_4k.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/a_bench_ref.json 2> gpurun_out/a_bench_ref.err
tail -3 gpurun_out/a_tests.log; tail -2 gpurun_out/a_noalloc_alone.log; tail -1 gpurun_out/a_bench_1080p.err; tail -1 gpurun_out/a_bench_4k.err
