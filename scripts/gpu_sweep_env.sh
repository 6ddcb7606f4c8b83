#!/bin/bash
# On the B200 box: time the front-end step under different settings of the overlap tunables (environment variables read
# by svo_gpu_create).  Usage: scripts/gpu_sweep_env.sh <tag> "VAR=val VAR=val" ["VAR=val ..." ...]
TAG=${1:-sweepenv}; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
i=0
for cfg in "$@"; do
  i=$((i+1))
  env $cfg timeout 300 python bench.py --steps 60 --warmup 6 --no-cpu-baseline > $OUT/c$i.json 2> $OUT/c$i.err
  python - "$cfg" "$OUT/c$i.json" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print("%-70s step %.3f ms  klt %.3f  fast %.3f  pyr %.3f  bucket %.3f  e2e %.0f fps  ok %.6f" % (sys.argv[1], d["ms_per_step"],
          d["kernels"]["klt"]["ms"], d["kernels"]["fast"]["ms"], d["kernels"]["pyramid"]["ms"], d["kernels"]["bucket"]["ms"],
          d["e2e"]["value"], d["stats"]["status_ok_frac"]))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done | tee $OUT/summary.txt
