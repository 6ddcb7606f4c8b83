#!/bin/bash
# Quick GPU loop: LK/step parity tests + one short bench line.  Usage: scripts/gpu_quick.sh <tag> [pytest -k expr]
set -u
TAG=${1:-q}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q ${2:+-k "$2"} > $OUT/pytest.log 2>&1; echo "pytest rc=$?"
tail -15 $OUT/pytest.log
timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-extras > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
python - <<PY
import json
d = json.load(open("$OUT/bench.json"))
print("value %.0f e2e %.0f ms/step %.4f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]))
print({k: round(v["ms"], 4) for k, v in d["kernels"].items() if isinstance(v, dict)})
PY
