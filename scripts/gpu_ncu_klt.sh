#!/bin/bash
# One ncu --set full capture of the LK kernel only (after a plain run exited 0).  Usage: scripts/gpu_ncu_klt.sh <tag>
set -u
TAG=${1:-k}
OUT=gpurun_out/$TAG
mkdir -p $OUT
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 600 $SHORT > $OUT/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:klt_' -s 4 -c 1 -o $OUT/klt $SHORT > $OUT/ncu.log 2>&1
echo "ncu rc=$?"
