#!/bin/bash
# one stream: FAST CTA size and bucketing formulation.  scripts/gpu_single_sweep.sh <tag>
OUT=gpurun_out/${1:-single}; mkdir -p $OUT
for spec in "1 0" "0 0" "2 0" "1 2" "2 2"; do
  set -- $spec
  SVO_GPU_FAST_SMALL=$1 SVO_GPU_BUCKET_MODE=$2 timeout 600 python bench.py --streams 1 --steps 300 --warmup 10 --no-cpu-baseline --no-extras > $OUT/f$1_b$2.json 2> $OUT/f$1_b$2.err
  python - <<PY
import json
d = json.load(open("$OUT/f$1_b$2.json")); k = d["kernels"]
print("fast_small=$1 bucket_mode=$2: %.4f ms/frame dev, e2e %.4f | pyr %.4f klt %.4f fast %.4f bucket %.4f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], k["pyramid"]["ms"], k["klt"]["ms"], k["fast"]["ms"], k["bucket"]["ms"]))
PY
done
