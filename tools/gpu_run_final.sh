#!/bin/bash
# Final single-GPU call of a round: whole GPU suite, the bench lines that go to profiles/, ncu launch list and full
# captures of the search kernels. Usage: gpu_run_final.sh TAG
TAG=${1:-r2}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log; tail -2 gpurun_out/${TAG}_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -1 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference_arm.json 2> gpurun_out/${TAG}_ref.err
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_1080p_p8.json 2> gpurun_out/${TAG}_1080p.err; echo "1080p rc=$?"
timeout 900 python bench.py --workload 4k_p4 --steps 4 --warmup 3 > gpurun_out/${TAG}_bench_4k_p4.json 2> gpurun_out/${TAG}_4k.err; echo "4k rc=$?"
timeout 900 python bench.py --workload 8k_p6 --steps 4 --warmup 3 > gpurun_out/${TAG}_bench_8k_p6_1gpu.json 2> gpurun_out/${TAG}_8k.err; echo "8k rc=$?"
timeout 300 python bench.py --workload kernel_1080p --steps 10 --warmup 3 --cpu-seconds 5 > gpurun_out/${TAG}_bench_kernel_1080p.json 2> gpurun_out/${TAG}_k.err; echo "kernel rc=$?"
for wl in 1080p_p8 4k_p4; do timeout 300 python tools/prof_stats.py $wl 5 2>&1 | tail -1; done > gpurun_out/${TAG}_stats_kernels.jsonl
for wl in 1080p_p8 4k_p4; do timeout 300 python tools/prof_dg.py $wl 6 2>&1 | tail -1; done > gpurun_out/${TAG}_dg_kernel.jsonl
CMD="python bench.py --steps 2 --warmup 3 --cpu-seconds 0.5 --cpu-pictures 2"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
python tools/prof_phase.py 1080p_p8 1 > gpurun_out/${TAG}_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"strip_kernel|int_kernel|refine_kernel|finish_kernel|points_kernel|scalar_kernel" -s 22 -c 11 -o gpurun_out/${TAG}_prof_stage python tools/prof_phase.py 1080p_p8 1 > gpurun_out/${TAG}_ncu_full.log 2>&1; tail -2 gpurun_out/${TAG}_ncu_full.log
python tools/bench_summary.py gpurun_out/${TAG}_bench_1080p_p8.json gpurun_out/${TAG}_bench_4k_p4.json gpurun_out/${TAG}_bench_8k_p6_1gpu.json
