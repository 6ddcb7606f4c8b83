#!/bin/bash
# one 8-GPU box: weak scaling N = 1, 2, 4, 8 (64 streams per GPU), strong scaling (64 streams in total, BASELINE config 5 as
# written) N = 2, 4, 8, and the CPU arm.  Usage: scripts/gpu_scale.sh <tag>
OUT=gpurun_out/${1:-scale}; mkdir -p $OUT
nvidia-smi topo -m > $OUT/topo.txt 2>&1
lscpu | head -25 > $OUT/lscpu.txt 2>&1
numactl -H > $OUT/numa.txt 2>&1 || ls /sys/devices/system/node > $OUT/numa.txt
run() {  # n extra-args out
  if [ "$1" = 1 ]; then timeout 600 python bench.py --gpus 1 --steps 200 --warmup 10 --no-cpu-baseline $2 > $OUT/$3.json 2> $OUT/$3.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29500 + $1)) bench.py --gpus $1 --steps 200 --warmup 10 --no-cpu-baseline $2 > $OUT/$3.json 2> $OUT/$3.err; fi
  python - <<PY
import json
try:
    d = [json.loads(l) for l in open("$OUT/$3.json") if l.startswith("{")][-1]
    print("$3: n=%d value %.0f e2e %.0f ms/step %.4f e2e_ms %.4f h2d %s" % (d["n_gpus"], d["value"], d["e2e"]["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], json.dumps(d.get("h2d_probe", {}).get("gb_per_s_per_gpu"))))
except Exception as e:
    print("$3: failed", e)
PY
}
for n in 1 2 4 8; do run $n "" weak$n; done
for n in 2 4 8; do run $n "--total-streams 64" strong$n; done
timeout 600 python bench.py --impl reference --gpus 8 --steps 20 --warmup 5 > $OUT/reference.json 2> $OUT/reference.err; tail -c 600 $OUT/reference.json
