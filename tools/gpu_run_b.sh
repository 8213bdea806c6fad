#!/bin/bash
# GPU call B: parity tests (stage-level) + bench at 1080p and 4K. Usage: gpu_run_b.sh TAG [pytest -k expr]
TAG=${1:-b}; K=${2:-""}
mkdir -p gpurun_out
if [ -n "$K" ]; then KARG=(-k "$K"); else KARG=(); fi
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_big.py tests/test_encoder_dump.py -m gpu -x -q "${KARG[@]}" > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 --cpu-seconds 2 > gpurun_out/${TAG}_bench_1080p.json 2> gpurun_out/${TAG}_bench_1080p.err; echo "rc=$?" >> gpurun_out/${TAG}_bench_1080p.err
timeout 900 python bench.py --workload 4k_p4 --steps 3 --warmup 3 --cpu-seconds 2 > gpurun_out/${TAG}_bench_4k.json 2> gpurun_out/${TAG}_bench_4k.err; echo "rc=$?" >> gpurun_out/${TAG}_bench_4k.err
tail -4 gpurun_out/${TAG}_tests.log; tail -2 gpurun_out/${TAG}_bench_1080p.err; tail -2 gpurun_out/${TAG}_bench_4k.err
python tools/bench_summary.py gpurun_out/${TAG}_bench_1080p.json gpurun_out/${TAG}_bench_4k.json
