#!/bin/bash
# multi-GPU call: gpu_run_mg.sh N TAG [runs...] — picture-sharded / stream-sharded benches on N GPUs
N=${1:-2}; TAG=${2:-mg}; shift 2; RUNS=${@:-"4k_pic 8k_pic 1080p_stream"}
mkdir -p gpurun_out
run() { # name, args...
  local name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@" \
     > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err; echo "$name rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*" gpurun_out/${TAG}_${name}.err | tail -3 | cut -c1-300
}
for r in $RUNS; do
  case $r in
    1080p_pic) run 1080p_pic --workload 1080p_p8 --sharding picture --pyramids nccl --steps 4 --warmup 3 --cpu-seconds 2 ;;
    4k_pic) run 4k_pic --workload 4k_p4 --sharding picture --pyramids nccl --steps 4 --warmup 3 --cpu-seconds 2 ;;
    8k_pic) run 8k_pic --workload 8k_p6 --sharding picture --pyramids nccl --steps 4 --warmup 3 --cpu-seconds 2 ;;
    1080p_stream) run 1080p_stream --workload 1080p_p8 --steps 6 --warmup 3 --cpu-seconds 2 ;;
    4k_stream) run 4k_stream --workload 4k_p4 --steps 4 --warmup 3 --cpu-seconds 2 ;;
  esac
done
python tools/bench_summary.py gpurun_out/${TAG}_*.json
