#!/bin/bash
OUT=gpurun_out/${1:-ovl64}; mkdir -p $OUT
for o in 0 1 0 1; do
  SVO_GPU_STEP_OVERLAP=$o python bench.py --steps 300 --warmup 10 --no-cpu-baseline --no-extras > $OUT/o$o.json 2> $OUT/o$o.err
  python - <<PY
import json
d = json.load(open("$OUT/o$o.json"))
print("overlap=$o 64 streams: dev %.4f ms/step (%.0f frames/s) e2e %.4f (%.0f)" % (d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["e2e"]["value"]))
PY
done
