#!/bin/bash
# ncu launch list (gpu__time_duration per kernel) of a short bench run.  Usage: scripts/gpu_launches.sh <tag>
TAG=${1:-l}
OUT=gpurun_out/$TAG
mkdir -p $OUT
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 600 $SHORT > $OUT/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -c 600 --csv --log-file $OUT/launches.csv $SHORT > $OUT/ncu_launch.log 2>&1
echo "rc=$?"
