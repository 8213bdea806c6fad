#!/bin/bash
# multi-GPU call: gpu_run_mg.sh N TAG — sharing test + picture-sharded / stream-sharded benches on N GPUs
N=${1:-2}; TAG=${2:-mg}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "share_picture or two_handles" > gpurun_out/${TAG}_tests.log 2>&1; tail -2 gpurun_out/${TAG}_tests.log
run() { # name, args...
  local name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@" \
     > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err; echo "$name rc=$?"; tail -3 gpurun_out/${TAG}_${name}.err | cut -c1-300
}
run 1080p_pic --workload 1080p_p8 --sharding picture --pyramids nccl --steps 3 --warmup 3 --cpu-seconds 2
run 4k_pic --workload 4k_p4 --sharding picture --pyramids nccl --steps 3 --warmup 3 --cpu-seconds 2
if [ "$N" -ge 4 ]; then run 8k_pic --workload 8k_p6 --sharding picture --pyramids nccl --steps 3 --warmup 3 --cpu-seconds 2; fi
run 1080p_stream --workload 1080p_p8 --steps 5 --warmup 3 --cpu-seconds 2
python tools/bench_summary.py gpurun_out/${TAG}_*.json
