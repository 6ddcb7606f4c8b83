#!/bin/bash
# the other BASELINE configs as bench lines (device-resident + e2e), plus a probe for C++ OpenCV on the box
OUT=gpurun_out/${1:-cfg}; mkdir -p $OUT
( pkg-config --cflags opencv4; find / -name "core.hpp" -path "*opencv2*" 2>/dev/null | head -5; find / -name "libopencv_core*" 2>/dev/null | head -5 ) > $OUT/opencv_probe.txt 2>&1
cat $OUT/opencv_probe.txt
for spec in "C2 32" "C3 8" "C4 2" "C1w11 64" "C1 1" "C1 4" "C1 16"; do
  set -- $spec
  timeout 600 python bench.py --workload $1 --streams $2 --steps 100 --warmup 5 --no-cpu-baseline --no-extras > $OUT/$1_$2.json 2> $OUT/$1_$2.err
  python - <<PY
import json
try:
    d = json.load(open("$OUT/$1_$2.json"))
    k = d["kernels"]
    print("%-6s streams %3d: %9.0f frames/s dev, %9.0f e2e, %.4f ms/step | pyr %.4f klt %.4f fast %.4f bucket %.4f | tracks/s %.3g kp %.0f" % ("$1", $2, d["value"], d["e2e"]["value"], d["ms_per_step"], k["pyramid"]["ms"], k["klt"]["ms"], k["fast"]["ms"], k["bucket"]["ms"], d["klt_tracks_per_s"], d["stats"]["keypoints_per_frame"]))
except Exception as e:
    print("$1 $2 failed", e, open("$OUT/$1_$2.err").read()[-300:])
PY
done
