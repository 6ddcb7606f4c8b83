"""The N>1 path on CPU: world_size-2 gloo process group exercising the sequence sharding, the max-over-ranks timing
reduction and bench.py's rank-0-only reference arm (replicas only, no data-path collective)."""
import json
import os
import subprocess
import sys

import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from svo_b200 import sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_properties():
    for world in (1, 2, 4, 8):
        owned = [sharding.sequences_of(r, world, 64) for r in range(world)]
        assert sorted(sum(owned, [])) == list(range(64))                       # every sequence exactly once
        assert max(map(len, owned)) - min(map(len, owned)) == 0                  # C5's 64 sequences split evenly
        seeds = [sharding.stream_seeds(r, 64, 11000) for r in range(world)]
        flat = sum(seeds, [])
        assert len(set(flat)) == len(flat) and seeds[0] == sharding.stream_seeds(0, 64, 11000)
    assert sharding.sequences_of(1, 3, 7) == [1, 4]                              # ragged split
    assert sharding.aggregate_rate(64 * 10, 2, 20.0) == pytest.approx(64000.0)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = sharding.sequences_of(rank, world, 7)
        ms = sharding.max_over_ranks(10.0 + 5.0 * rank, world)                   # rank 1 is the slow one
        dist.barrier()
        q.put((rank, mine, ms))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_max_over_ranks_and_partition():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert out[0][1] == [0, 2, 4, 6] and out[1][1] == [1, 3, 5]
    assert out[0][2] == out[1][2] == 15.0                                        # both ranks see the slowest rank's time


def test_reference_arm_prints_on_rank0_only():
    """bench.py --impl reference under a 2-rank launch: rank 0 runs and prints the CPU line, rank 1 exits 0 silently."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + os.getpid() % 300), os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
           "--steps", "1", "--warmup", "1", "--streams", "1", "--workload", "C2", "--cpu-single-frames", "2"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["value"] > 0 and d["gpu_launches"] == 0
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["cpu_baseline"]["kind"] in ("reference", "port")
    cb = d["cpu_baseline"]
    assert cb["python_overhead_frac"] < 0.02 and cb["single_sequence_1_thread"]["frames_per_s"] > 0
    assert cb["cv2"]["getNumberOfCPUs"] >= 1 and set(cb["stage_ms_median"]) == {"lk", "mask", "detect", "convert", "bucket"}
