#!/bin/bash
# On the B200 box: time the front-end step with every tuning variant under svo_b200/lib/variants/ (plus the default build).
# Usage: scripts/gpu_sweep.sh <tag>
TAG=${1:-sweep}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for so in svo_b200/lib/libsvo_gpu.so svo_b200/lib/variants/*.so; do
  name=$(basename $so .so)
  SVO_GPU_LIB=$PWD/$so timeout 300 python bench.py --steps 30 --warmup 6 --no-cpu-baseline > $OUT/$name.json 2> $OUT/$name.err
  python - "$name" "$OUT/$name.json" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print("%-12s step %.3f ms  klt %.3f  fast %.3f  pyr %.3f  bucket %.3f  e2e %.0f fps  ok %.6f" % (sys.argv[1], d["ms_per_step"],
          d["kernels"]["klt"]["ms"], d["kernels"]["fast"]["ms"], d["kernels"]["pyramid"]["ms"], d["kernels"]["bucket"]["ms"],
          d["e2e"]["value"], d["stats"]["status_ok_frac"]))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done | tee $OUT/summary.txt
