#!/bin/bash
# the two ends of the scaling table on one 8-GPU box: weak N = 8 (64 streams per GPU) and strong N = 8 (64 streams in total)
OUT=gpurun_out/${1:-scale8}; mkdir -p $OUT
run() {
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) bench.py --gpus 8 --steps 200 --warmup 10 --no-cpu-baseline $1 > $OUT/$2.json 2> $OUT/$2.err
  python - <<PY
import json
try:
    d = [json.loads(l) for l in open("$OUT/$2.json") if l.startswith("{")][-1]
    print("$2: n=%d value %.0f e2e %.0f ms/step %.4f e2e_ms %.4f" % (d["n_gpus"], d["value"], d["e2e"]["value"], d["ms_per_step"], d["e2e"]["ms_per_step"]))
except Exception as e:
    print("$2: failed", e)
PY
}
run "" weak8
run "--total-streams 64" strong8
