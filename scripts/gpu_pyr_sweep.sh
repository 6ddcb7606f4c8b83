#!/bin/bash
# fused vs per-level pyramid at 64 streams and at 1 stream: scripts/gpu_pyr_sweep.sh <tag>
OUT=gpurun_out/${1:-pyr}; mkdir -p $OUT
for f in 0 1; do
  SVO_GPU_PYR_FUSED=$f timeout 600 python bench.py --steps 200 --warmup 10 --no-cpu-baseline > $OUT/fused$f.json 2> $OUT/fused$f.err
  python - <<PY
import json
d = json.load(open("$OUT/fused$f.json"))
s = d["single_stream"]
print("fused=$f value %.0f e2e %.0f ms/step %.4f pyramid %.4f | single dev %.4f e2e %.4f sync %.4f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["kernels"]["pyramid"]["ms"], s["device_resident"]["ms_per_frame"], s["e2e"]["ms_per_frame"], s["e2e_synchronous"]["ms_per_frame"]))
PY
done
