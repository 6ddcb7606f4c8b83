#!/bin/bash
# Run on the B200 box through gpurun: GPU parity tests, one bench line, the ncu launch list and one full capture.
# Usage: scripts/gpu_check.sh <tag> [skip_tests]
set -u
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
if [ "${2:-}" != "skip_tests" ]; then
  timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest.log
  tail -5 $OUT/pytest.log
fi
timeout 900 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
cat $OUT/bench.json
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
timeout 600 $SHORT > $OUT/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches.csv $SHORT > $OUT/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
timeout 600 $SHORT > $OUT/plain2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k 'regex:klt_|fast_|pyrdown|ingest|bucket_|seed_tracks|mask_boxes' -s 65 -c 12 -o $OUT/prof $SHORT > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT
