#!/bin/bash
# bench_variants.sh NAME... : time each experimental library of build_alt/ with bench.py (GPU box)
cd "$(dirname "$0")/.."
L=svt-av1-mod-by-patman_b200/csrc/libsvtme_b200.so
cp $L /tmp/libsvtme_keep.so
for v in "$@"; do
  cp build_alt/libsvtme_$v.so $L
  python bench.py --steps 3 --warmup 3 --cpu-seconds 0.3 ${BENCH_ARGS} > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err
  python - "$v" <<'PY'
import json,sys
v=sys.argv[1]
try:
    d=json.load(open("gpurun_out/var_%s.json"%v)); r=d["roofline"]
    print("%-10s %.3f ms  %s  frac %.4f pipe %.4f"%(v,d["ms_per_step"],{k:x for k,x in r["phase_ms"].items() if x>0.2},r["frac"],r["pipeline"]["frac"]))
except Exception as e:
    print(v,"FAILED",e)
PY
done
cp /tmp/libsvtme_keep.so $L
