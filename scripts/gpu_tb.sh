#!/bin/bash
# GPU box: all GPU tests + one full bench line (extras and CPU baseline included).  Usage: scripts/gpu_tb.sh <tag> [pytest -k expr] [bench args]
set -u
TAG=${1:-t}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q ${2:+-k "$2"} > $OUT/pytest.log 2>&1; echo "pytest rc=$?"
tail -15 $OUT/pytest.log
timeout 900 python bench.py ${3:-} > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
tail -5 $OUT/bench.err
python - <<PY
import json
d = json.load(open("$OUT/bench.json"))
print("value %.0f e2e %.0f ms/step %.4f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]))
print({k: round(v["ms"], 4) for k, v in d["kernels"].items() if isinstance(v, dict)})
print("single", json.dumps(d.get("single_stream")))
print("drop_in", json.dumps(d.get("drop_in")))
cb = d.get("cpu_baseline") or {}
print("cpu", cb.get("value"), cb.get("cores"))
PY
