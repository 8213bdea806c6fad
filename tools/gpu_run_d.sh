#!/bin/bash
# GPU call D: end-to-end schedule of the 1080p workload for several batch sizes (pictures per submitted batch)
mkdir -p gpurun_out
for ch in "$@"; do
  tag=$(echo $ch | tr ',' '_')
  timeout 300 python bench.py --steps ${STEPS:-6} --warmup 3 --cpu-seconds 2 --e2e-chunks $ch > gpurun_out/d_$tag.json 2> gpurun_out/d_$tag.err
  python - <<PY
import json
r=json.load(open("gpurun_out/d_$tag.json"))
e=r["e2e"]
print("chunks $ch: resident %.3f ms, e2e best %.3f (%s), streamed %.3f, drained %.3f, roofline frac %.3f slot %.3f" % (r["ms_per_step"], e["ms_per_step"], e["schedule"][:8], e["streamed_ms_per_step"], e["drained"]["ms_per_step"], r["roofline"]["frac"], r["roofline"]["slot"]["frac"]))
PY
done
