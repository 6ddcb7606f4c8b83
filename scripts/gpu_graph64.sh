#!/bin/bash
OUT=gpurun_out/${1:-g64}; mkdir -p $OUT
for g in 0 1 0 1; do
  SVO_GPU_STEP_GRAPH=$g python bench.py --steps 300 --warmup 10 --no-cpu-baseline --no-extras > $OUT/g$g.json 2> $OUT/g$g.err
  python - <<PY
import json
d = json.load(open("$OUT/g$g.json"))
print("graph=$g 64 streams: dev %.4f ms/step (%.0f frames/s) e2e %.4f (%.0f)" % (d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["e2e"]["value"]))
PY
done
for s in 2 4 8 16; do
  for g in 0 1; do
  SVO_GPU_STEP_GRAPH=$g python bench.py --streams $s --steps 300 --warmup 10 --no-cpu-baseline --no-extras > $OUT/s${s}g$g.json 2> $OUT/s${s}g$g.err
  python - <<PY
import json
d = json.load(open("$OUT/s${s}g$g.json"))
print("graph=$g $s streams: dev %.4f ms/step (%.0f frames/s) e2e %.4f (%.0f)" % (d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["e2e"]["value"]))
PY
  done
done
