#!/bin/bash
# On the B200 box: one bench line per BASELINE.json config shape that is not the headline (C2, C3, C4), device-resident
# + end to end, no CPU baseline.  Usage: scripts/gpu_configs.sh <tag>
TAG=${1:-cfgs}
OUT=gpurun_out/$TAG
mkdir -p $OUT
run() {  # workload streams steps
  timeout 600 python bench.py --workload $1 --streams $2 --steps $3 --warmup 4 --no-cpu-baseline > $OUT/$1.json 2> $OUT/$1.err
  python - "$1" "$OUT/$1.json" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print("%s streams %d: %.0f frames/s (e2e %.0f), %.3g tracks/s, step %.3f ms, stages %s, it/track %.1f, ok %.5f" % (
        sys.argv[1], d["config"]["streams_per_gpu"], d["value"], d["e2e"]["value"], d["klt_tracks_per_s"], d["ms_per_step"],
        {k: round(v["ms"], 3) for k, v in d["kernels"].items() if isinstance(v, dict)}, d["klt_issue_model"]["iterations_per_track"],
        d["stats"]["status_ok_frac"]))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
{ run C2 64 100; run C3 16 60; run C4 4 30; } | tee $OUT/summary.txt
