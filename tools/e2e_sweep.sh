#!/bin/bash
# e2e_sweep.sh : end-to-end step time (tools/e2e_timeline.py) over compute-stream counts and chunkings (GPU box)
cd "$(dirname "$0")/.."
for ns in 1 2 3 4; do
  for cfg in "4,12,16 2" "4,8 3" "4,8,12,16 3" "8 4"; do
    set -- $cfg
    SVTME_COMPUTE_STREAMS=$ns python tools/e2e_timeline.py $1 $2 > gpurun_out/tl.txt 2>&1
    echo "streams $ns chunks $1 depth $2: $(head -1 gpurun_out/tl.txt | grep -o 'collect[0-9]*@[0-9.]*' | tail -1)"
  done
done
