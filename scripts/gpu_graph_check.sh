#!/bin/bash
# GPU tests, host submit cost, one stream with and without the captured step graph, one full bench line
OUT=gpurun_out/${1:-graph}; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python scripts/gpu_submit_cost.py
for g in 0 1; do
  SVO_GPU_STEP_GRAPH=$g python bench.py --streams 1 --steps 300 --warmup 10 --no-cpu-baseline --no-extras > $OUT/g$g.json 2> $OUT/g$g.err
  python - <<PY
import json
d = json.load(open("$OUT/g$g.json"))
print("graph=$g 1 stream: dev %.4f ms/frame e2e %.4f launches %d" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["gpu_launches"]))
PY
done
python bench.py --steps 100 --warmup 10 --no-cpu-baseline > $OUT/full.json 2> $OUT/full.err
python - <<PY
import json
d = json.load(open("$OUT/full.json"))
print(d["value"], d["e2e"]["value"])
print(json.dumps(d["single_stream"]))
print(json.dumps(d["roofline"].get("issue_slots")))
print(json.dumps(d["drop_in"]))
PY
