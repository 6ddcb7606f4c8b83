#!/bin/bash
# detection beside LK (SVO_GPU_STEP_OVERLAP) forced off / on at several stream counts
OUT=gpurun_out/${1:-ovl}; mkdir -p $OUT
for s in 1 2 4 8 16 32; do
  for o in 0 1; do
  SVO_GPU_STEP_OVERLAP=$o python bench.py --streams $s --steps 300 --warmup 10 --no-cpu-baseline --no-extras > $OUT/s${s}o$o.json 2> $OUT/s${s}o$o.err
  python - <<PY
import json
d = json.load(open("$OUT/s${s}o$o.json"))
print("overlap=$o $s streams: dev %.4f ms/step (%.0f frames/s) e2e %.4f (%.0f)" % (d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["e2e"]["value"]))
PY
  done
done
