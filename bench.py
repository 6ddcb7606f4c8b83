#!/usr/bin/env python
"""bench.py -- front-end frames/s of the svo per-frame feature front-end (FAST + bucketing + 4-level 21x21 KLT).

One "step" = one frame of every stream on this GPU through svo_gpu_step (libsvo_gpu.so, C ABI): ingest + pyramid,
LK from the previous frame on each stream's 2k tracks, the +-10 px mask around those tracks, FAST-9/16 + NMS on the new
frame, grid bucketing, re-seeding of the tracks (src/tracking.cpp:240-257 on a key-frame).  Workload = BASELINE.json's
metric configuration (1241x376 KITTI-shaped synthetic frames, 2k tracks per stream) with `--streams` independent sequences
per GPU (C5: 64).  Streams never interact, so N GPUs = N x streams sequences with no collective ("scaling": "weak");
`--total-streams T` instead shards a FIXED set of T sequences over the ranks (BASELINE config 5 as written, "strong").
torch.distributed is used for the start/stop barriers only.

  value : frames/s with the frames already resident in HBM (device ring larger than L2), CUDA events on the context's stream
  e2e   : frames/s through the same C-ABI call with HOST frames (pinned), H2D of every frame and D2H of the step's
          results (tracked points, status, err, bucketed features, counts) inside the timed region
  single_stream : BASELINE config 1 (ONE sequence): ms/frame device-resident and end to end through the step API
  drop_in       : config 1 through the reference-signature C++ wrappers (GpuFastDetector::detect ->
                  FeatureSet::bucketingFeatures -> svo::calcOpticalFlowPyrLK on host cv::Mat frames)
  --impl reference : the reference's own CPU path (OpenCV via cv2 + src/bucket.cpp compiled unmodified) on the host cores,
                     one process per core (oracle/cpu_arm.py)
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from svo_b200 import sharding, synth  # noqa: E402

RING = 8   # frames per stream kept resident; time stamps 0,1,2,3,4,3,2,1 so that consecutive frames (and the wrap) are ~2 px apart
METRIC = "front-end frames/s (FAST+bucketing+4-level 21x21 KLT, 2k tracks/stream)"
DTYPE = "u8/int32 + f32 epilogue"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--streams", type=int, default=64, help="independent sequences per GPU (weak scaling)")
    ap.add_argument("--total-streams", type=int, default=0,
                    help="strong scaling: a fixed set of this many sequences sharded over the ranks (config 5: 64)")
    ap.add_argument("--workload", default="C1", choices=sorted(synth.WORKLOADS), help="per-stream shape")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the single_stream / drop_in / h2d_probe blocks")
    ap.add_argument("--cpu-frames", type=int, default=12, help="frames per stream of the cpu_baseline aggregate sample")
    ap.add_argument("--cpu-single-frames", type=int, default=50, help="frames of the single-sequence CPU figures (T = 1, T = all)")
    return ap.parse_args()


def stream_ids(args, rank, world):
    """Global sequence ids this rank owns."""
    if args.total_streams:
        return sharding.sequences_of(rank, world, args.total_streams)
    return list(range(rank * args.streams, (rank + 1) * args.streams))


def make_frames(wl, ids, seed0):
    """uint8 [RING, len(ids), h, w]: stream j = synthetic sequence with seed seed0 + ids[j]."""
    w, h = wl["w"], wl["h"]
    times = synth.ring_times(RING)
    distinct = sorted(set(times))

    def one(g):
        canvas = synth.make_canvas(w, h, seed0 + g)
        fr = {t: synth.render_frame(canvas, w, h, t) for t in distinct}
        return np.stack([fr[t] for t in times])

    with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as ex:
        seqs = list(ex.map(one, ids))
    return np.ascontiguousarray(np.stack(seqs, axis=1))


def step_params(wl):
    return dict(fast_threshold=wl["fast_thr"], bucket_size=wl["bucket_size"], per_bucket=wl["per_bucket"],
                max_tracks=wl["n_tracks"], win_w=wl["win"], win_h=wl["win"], max_level=wl["max_level"], mask_half=10.0)


def cpu_params(wl):
    from oracle.cpu_pipeline import default_params
    return default_params(wl["w"], wl["h"], wl["fast_thr"], wl["bucket_size"], wl["per_bucket"], wl["n_tracks"], wl["win"],
                          wl["max_level"], mask_half=10.0)


def pyramid_pixels(w, h, levels):
    n = 0
    for _ in range(levels):
        n += w * h
        w, h = (w + 1) // 2, (h + 1) // 2
    return n


class ClockSampler:
    """SM clock, power and throttle reasons during the timed region (B200_PROFILING.md recipe), read in-process through NVML:
    a thread polls every 2 ms from before the warm-up, and mark() takes one sample inline -- called right after the timed
    region starts and right before it is synchronised, so even a 25 ms region (--steps 20) has samples.  An nvidia-smi
    subprocess needs longer than that to produce its first line."""
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, index):
        import threading
        self.rows, self.h, self.nv, self.err = [], None, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if ids and index < len(ids):
                self.h = (pynvml.nvmlDeviceGetHandleByIndex(int(ids[index])) if ids[index].isdigit()
                          else pynvml.nvmlDeviceGetHandleByUUID(ids[index]))
            else:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nv = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:   # no NVML: the line says so instead of inventing clocks
            self.err = "%s: %s" % (type(e).__name__, e)
            self.h = None
        self.lock = threading.Lock()
        self.run = self.h is not None
        self.t = threading.Thread(target=self._poll, daemon=True) if self.run else None
        if self.t:
            self.t.start()

    def _sample(self):
        nv = self.nv
        mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
        try:
            watts = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
        except Exception:
            watts = float("nan")
        try:
            get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
            mask = int(get(self.h))
        except Exception:
            mask = 0
        with self.lock:
            self.rows.append((time.time(), mhz, watts, mask))

    def _poll(self):
        while self.run:
            try:
                self._sample()
            except Exception:
                pass
            time.sleep(0.002)

    def mark(self):
        if self.h is not None:
            try:
                self._sample()
            except Exception:
                pass

    def stop(self, t0=None, t1=None):
        """Summary of the samples taken between host times t0 and t1 (time.time()); all samples if none fall inside."""
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": "nvml (in-process, 2 ms poll + 2 inline samples)"}
        if self.h is None:
            out["source"] = "unavailable (%s)" % self.err
            return out
        self.run = False
        self.t.join(timeout=2)
        with self.lock:
            rows = list(self.rows)
        inside = [r for r in rows if t0 is not None and t0 <= r[0] <= t1]
        use = inside or rows
        if use:
            mask = 0
            for r in use:
                mask |= r[3]
            out.update(sm_mhz=float(np.median([r[1] for r in use])), sm_max_mhz=self.max_mhz,
                       power_w=float(np.nanmedian([r[2] for r in use])), reasons=[n for b, n in self.REASONS if mask & b],
                       samples=len(use), window="timed region" if inside else "whole run (no sample fell inside the timed region)")
        return out


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def workload_config(args, wl, world):
    """Identical for both arms (the CPU arm's sample is described in its cpu_baseline block)."""
    if args.total_streams:
        what = "%d independent %dx%d synthetic sequences sharded over %d GPU(s)" % (args.total_streams, wl["w"], wl["h"], world)
    else:
        what = "%d independent %dx%d synthetic sequences per GPU" % (args.streams, wl["w"], wl["h"])
    return {"workload": "C5-shaped: %s, %d tracks each (C1 per stream: mask +-10 px, FAST thr %d + NMS, bucketing %d px / %d "
                        "per cell, KLT %dx%d, %d levels, 50 it / eps 0.001, min-eig err)" %
                        (what, wl["n_tracks"], wl["fast_thr"], wl["bucket_size"], wl["per_bucket"], wl["win"], wl["win"],
                         wl["max_level"] + 1),
            "streams_per_gpu": args.streams if not args.total_streams else None,
            "total_streams": args.total_streams or None,
            "width": wl["w"], "height": wl["h"], "tracks_per_stream": wl["n_tracks"], "mask_half_px": 10,
            "parallelism": "independent sequences per GPU, no collective",
            "l2": "frame ring of %d frames per stream cycled (64 streams: 246 MB), larger than the 126 MB L2" % RING}


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_baseline_block(wl, warmup, steps, single_frames=50):
    """The reference's CPU path on this host (oracle/cpu_arm.py: processes, no per-keypoint Python) -> cpu_baseline dict."""
    from oracle import cpu_arm                       # bench.py's CPU arm is one of the places allowed to run oracle/
    rep = cpu_arm.full_report(wl, cpu_params(wl), wl["seed"] * 1000, warmup, steps, single_frames=single_frames)
    agg, t1, tall = rep["aggregate"], rep["t1"], rep["tall"]
    return {"value": agg["frames_per_s"], "unit": "frames/s", "cores": rep["cores"], "kind": agg["kind"],
            "sample": "%d processes (one sequence each, cv2.setNumThreads(1)) x %d timed frames after %d warm-up frames, common "
                      "start barrier; same per-stream workload as the GPU arm (mask on)" % (agg["processes"], steps, warmup),
            "seconds": agg["seconds"], "python_overhead_frac": agg["python_overhead_frac"],
            "stage_ms_median": agg["stage_ms_median"],
            "single_sequence_1_thread": {"frames_per_s": t1["frames_per_s"], "ms_per_frame_median": t1["frame_ms_median"],
                                         "stage_ms_median": t1["stage_ms_median"], "frames": t1["frames"]},
            "single_sequence_all_threads": {"frames_per_s": tall["frames_per_s"], "ms_per_frame_median": tall["frame_ms_median"],
                                            "stage_ms_median": tall["stage_ms_median"], "frames": tall["frames"],
                                            "cv_threads": tall["cv_threads"],
                                            "note": "the reference's own configuration: one stream, OpenCV parallel_for_ inside LK"},
            "cv2": rep["cv2"]}


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    wl = synth.WORKLOADS[args.workload]
    cb = cpu_baseline_block(wl, args.warmup, args.steps, args.cpu_single_frames)
    fps = cb["value"]
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * cb["seconds"] / args.steps, "higher_is_better": True,
        "scaling": "strong" if args.total_streams else "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
        "config": workload_config(args, wl, max(world, args.gpus)),
        "cpu_baseline": cb,
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "notes": "CPU arm: a step = one frame of each of %d streams (one per host core); value = all frames / wall time "
                 "between the common start and the last process's finish" % cb["cores"],
    }
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ B200 arm
def bind_near_gpu(local):
    """Keeps this rank's host threads (and therefore its first-touch pinned pages) on the NUMA node of its GPU, when the box
    has more than one; returns a description for the bench line.  No-op on single-node hosts."""
    try:
        bus = subprocess.run(["nvidia-smi", "-i", os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",")[local] if
                              os.environ.get("CUDA_VISIBLE_DEVICES") else str(local), "--query-gpu=pci.bus_id",
                              "--format=csv,noheader"], capture_output=True, text=True, timeout=20).stdout.strip().lower()
        if len(bus.split(":")[0]) == 8:       # 00000000:1b:00.0 -> 0000:1b:00.0 (sysfs form)
            bus = bus[4:]
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read())
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()]
        if node < 0 or len(nodes) < 2:
            return {"numa_node": node, "numa_nodes": len(nodes), "bound": False}
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return {"numa_node": node, "numa_nodes": len(nodes), "bound": bool(cpus), "cpus": len(cpus)}
    except Exception as e:
        return {"bound": False, "error": repr(e)[:80]}


def source_sha():
    h = hashlib.sha1()
    for f in ("klt.cu", "klt_common.cuh"):
        h.update(open(os.path.join(ROOT, "svo_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]


def run_single_stream(capi, Frontend, torch, local, wl, frames1, steps):
    """BASELINE config 1: ONE sequence through the step API (1 lane), two steps outstanding.  -> dict."""
    w, h = wl["w"], wl["h"]
    pitch = (w + 127) // 128 * 128
    n = max(steps, 50)
    dev = torch.zeros((RING, 1, h, pitch), dtype=torch.uint8, device="cuda")
    dev[..., :w] = torch.from_numpy(frames1).cuda()
    host = torch.from_numpy(frames1).pin_memory()
    cells = (h // wl["bucket_size"] + 1) * (w // wl["bucket_size"] + 1) * wl["per_bucket"]
    out = {}
    with capi.Context(device=local, max_w=w, max_h=h, max_levels=wl["max_level"] + 1, max_points=32768, max_batch=1) as ctx:
        fe = Frontend(ctx, w, h, 1, bucket_capacity=cells, **step_params(wl))
        stream = torch.cuda.ExternalStream(ctx.stream)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

        def loop(i0, i1, submit, depth):
            for i in range(i0, i1):
                submit(i)
                if fe.outstanding == depth:
                    fe.wait()
            while fe.outstanding:
                fe.wait()

        def sub_dev(i):
            fe.submit(dev[i % RING].data_ptr(), None, capi.MEM_DEVICE, pitch, h * pitch, with_outputs=False)

        def sub_host(i):
            fe.submit(host[i % RING].data_ptr(), host[(i + 1) % RING].data_ptr(), capi.MEM_HOST, w, w * h)

        for name, sub, depth in (("device_resident", sub_dev, 2), ("e2e", sub_host, 2), ("e2e_synchronous", sub_host, 1)):
            fe.reset()
            loop(0, 10, sub, depth)
            torch.cuda.synchronize()
            l0 = ctx.launch_count
            t0 = time.perf_counter()
            ev[0].record(stream)
            loop(10, 10 + n, sub, depth)
            ev[1].record(stream)
            torch.cuda.synchronize()
            wall = time.perf_counter() - t0
            out[name] = {"ms_per_frame": ev[0].elapsed_time(ev[1]) / n, "wall_ms_per_frame": 1e3 * wall / n,
                         "frames_per_s": n / wall, "launches_per_frame": (ctx.launch_count - l0) / n}
        out["frames"] = n
        out["note"] = ("one 1241x376 sequence, 2k tracks, through svo_gpu_step_submit/_wait; e2e_synchronous = each frame's "
                       "results are on the host before the next frame is submitted (the reference's loop, src/tracking.cpp:240-268)")
    return out


def run_drop_in(wl, frames1, n_frames=200):
    """Config 1 through the C++ wrappers with the reference's signatures (svo_b200/host/frontend_selftest --bench): host
    cv::Mat frames and std::vector containers in, std::vector<cv::KeyPoint> / Point2f / status / err out, every call."""
    exe = os.path.join(ROOT, "svo_b200", "lib", "frontend_selftest")
    if not os.path.exists(exe):
        return {"unavailable": "svo_b200/lib/frontend_selftest not built"}
    path = None
    try:
        with tempfile.NamedTemporaryFile("wb", suffix=".bin", delete=False) as f:
            path = f.name
            np.array([wl["w"], wl["h"], frames1.shape[0], wl["n_tracks"], wl["bucket_size"], wl["per_bucket"], wl["fast_thr"]],
                     np.int32).tofile(f)
            np.ascontiguousarray(frames1).tofile(f)
        r = subprocess.run([exe, "--bench", path, str(n_frames)], capture_output=True, text=True, timeout=300)
        for line in r.stdout.splitlines():
            if line.startswith("{"):
                return json.loads(line)
        return {"unavailable": "no JSON line (rc %d): %s" % (r.returncode, (r.stdout + r.stderr)[-200:])}
    except Exception as e:
        return {"unavailable": repr(e)[:200]}
    finally:
        if path and os.path.exists(path):
            os.unlink(path)


def run_b200(args):
    import torch
    import torch.distributed as dist
    from svo_b200 import capi
    from svo_b200.frontend import Frontend

    rank, local, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the front-end has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    binding = bind_near_gpu(local)
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's version banner (NCCL_DEBUG=VERSION on some boxes) goes to stderr
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "WARN"):
            os.environ.pop("NCCL_DEBUG")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = synth.WORKLOADS[args.workload]
    ids = stream_ids(args, rank, world)
    w, h, S, K, W = wl["w"], wl["h"], len(ids), args.steps, args.warmup
    total_streams = args.total_streams or world * args.streams
    levels = wl["max_level"] + 1
    pitch = (w + 127) // 128 * 128

    frames = make_frames(wl, ids, wl["seed"] * 1000)                     # [RING, S, h, w]
    dev = torch.zeros((RING, S, h, pitch), dtype=torch.uint8, device="cuda")
    dev[..., :w] = torch.from_numpy(frames).cuda()
    host = torch.from_numpy(frames).pin_memory()                         # compact rows, like a cv::Mat
    torch.cuda.synchronize()

    cells = (h // wl["bucket_size"] + 1) * (w // wl["bucket_size"] + 1) * wl["per_bucket"]
    ctx = capi.Context(device=local, max_w=w, max_h=h, max_levels=levels, max_points=max(32768, w * h // 16),
                       max_batch=S)
    fe = Frontend(ctx, w, h, S, bucket_capacity=cells, **step_params(wl))
    stream = torch.cuda.ExternalStream(ctx.stream)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(ms):
        return sharding.max_over_ranks(ms, world, device="cuda")

    # Both timed loops drive the step through svo_gpu_step_submit / svo_gpu_step_wait with two steps outstanding: step i+1 is
    # enqueued before the host waits for step i (frame i+1 never needs frame i's results on the HOST: the tracks are re-seeded
    # on the device), so the device does not idle while the host handles a frame.  Every step's results are waited for.
    def run_steps(i0, i1, submit, after_wait=None, depth=2, on_mid=None):
        for i in range(i0, i1):
            submit(i)
            if on_mid and i == (i0 + i1) // 2:
                on_mid()   # one clock sample with the loop half-way (about 0.1 ms of host time, the device has a step queued)
            if fe.outstanding == depth:
                fe.wait()
                if after_wait:
                    after_wait()
        while fe.outstanding:
            fe.wait()
            if after_wait:
                after_wait()

    # ---- end to end: host frames in, results out, every step.  Step i+1's frames (pinned host memory) are copied on a second
    #      stream while step i computes; every step's results land in pinned host arrays before its wait returns.
    def hptr(i):
        return host[i % RING].data_ptr()

    def submit_host(i):
        fe.submit(hptr(i), hptr(i + 1), capi.MEM_HOST, w, w * h)

    # Untimed steps in front of each timed loop: the W asked for, and at least 4, because the library records the step's chain
    # into CUDA graphs during steps 2 and 3 of a loop (six variants: 3 frame slots x 2 tickets); capture is set-up, not work.
    WU = max(W, 4)
    fe.reset()
    run_steps(0, WU, submit_host)
    barrier()
    ev[0].record(stream)
    run_steps(WU, WU + K, submit_host)
    ev[1].record(stream)
    barrier()
    e2e_ms = max_over_ranks(ev[0].elapsed_time(ev[1]))
    # statistics of the workload (outside any timed region)
    kp_sum = bk_sum = trk_sum = ok_sum = 0
    NS = 8
    for i in range(WU + K, WU + K + NS):
        fe.step_host_pipelined(hptr(i), hptr(i + 1), w, w * h)
        kp_sum += int(fe.n_keypoints.sum()); bk_sum += int(fe.n_bucketed.sum()); trk_sum += int(fe.n_tracked.sum())
        ok_sum += int(sum(int(fe.status[b, :fe.n_tracked[b]].sum()) for b in range(S)))

    # ---- H2D alone (same frames, same pinned ring, no kernels): names the limiter of e2e when ranks share a host
    h2d = None
    if not args.no_extras:
        scratch = torch.empty((S, h, w), dtype=torch.uint8, device="cuda")
        cs = torch.cuda.Stream()
        with torch.cuda.stream(cs):
            for i in range(3):
                scratch.copy_(host[i % RING], non_blocking=True)
            barrier()
            ev[4].record(cs)
            for i in range(40):
                scratch.copy_(host[i % RING], non_blocking=True)
            ev[5].record(cs)
        barrier()
        h2d_ms = max_over_ranks(ev[4].elapsed_time(ev[5])) / 40
        h2d = {"ms_per_step": h2d_ms, "gb_per_s_per_gpu": S * w * h / (h2d_ms * 1e-3) / 1e9,
               "gb_per_s_all_gpus": world * S * w * h / (h2d_ms * 1e-3) / 1e9,
               "note": "one step's frames (pinned host ring) copied alone, no kernels, all ranks at once, max over ranks"}
        del scratch

    # ---- device resident: frames already in HBM, nothing returned to the host
    sampler = ClockSampler(local) if rank == 0 else None
    def submit_dev(i):
        fe.submit(dev[i % RING].data_ptr(), None, capi.MEM_DEVICE, pitch, h * pitch, with_outputs=False)

    fe.reset()
    run_steps(0, WU, submit_dev)
    launches0 = ctx.launch_count
    barrier()
    t_start = time.time()
    ev[2].record(stream)
    acc = {"stage": np.zeros(4), "iters": 0, "n": 0}

    def after_wait():
        acc["stage"] += np.array(ctx.step_profile(True))
        acc["iters"] += ctx.last_klt_iterations
        acc["n"] += 1

    run_steps(WU, WU + K, submit_dev, on_mid=sampler.mark if sampler else None)
    ev[3].record(stream)
    if sampler:
        sampler.mark()   # the last steps are still running
    barrier()
    t_stop = time.time()
    dev_ms = max_over_ranks(ev[2].elapsed_time(ev[3]))
    launches = ctx.launch_count - launches0
    # per-stage times: the same steps one at a time (one step outstanding), so that no stage of step t+1 runs beside a stage
    # of step t and every CUDA-event bracket holds one stage alone.  Outside the timed region.
    ctx.step_profile(True)
    NP_ = 24
    run_steps(WU + K, WU + K + NP_, submit_dev, after_wait, depth=1)
    ctx.step_profile(False)
    clocks = sampler.stop(t_start, t_stop) if sampler else None
    stage, iters = acc["stage"], acc["iters"]

    # ---- config 5 as BASELINE.json words it: `--streams` sequences IN TOTAL, sharded over the ranks (strong split), measured
    #      by the same ranks right after the weak loops so that the driver's scaling run carries both.  Same step, same frames
    #      (the first len(ids) lanes of this rank's ring), a context of its own; outside the timed regions above.
    strong = None
    if world > 1 and not args.total_streams and not args.no_extras:
        tot = args.streams
        Ss = len(sharding.sequences_of(rank, world, tot))
        ctx_s = capi.Context(device=local, max_w=w, max_h=h, max_levels=levels, max_points=max(32768, w * h // 16), max_batch=Ss)
        fe_s = Frontend(ctx_s, w, h, Ss, bucket_capacity=cells, **step_params(wl))
        st_s = torch.cuda.ExternalStream(ctx_s.stream)
        es = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

        def loop_s(i0, i1, submit):
            for i in range(i0, i1):
                submit(i)
                if fe_s.outstanding == 2:
                    fe_s.wait()
            while fe_s.outstanding:
                fe_s.wait()

        def sub_h(i):
            fe_s.submit(hptr(i), hptr(i + 1), capi.MEM_HOST, w, w * h)

        def sub_d(i):
            fe_s.submit(dev[i % RING].data_ptr(), None, capi.MEM_DEVICE, pitch, h * pitch, with_outputs=False)

        strong = {"total_streams": tot, "streams_this_rank": Ss, "scaling": "strong", "steps": K, "warmup": WU}
        for key, sub in (("e2e", sub_h), ("value", sub_d)):
            fe_s.reset()
            loop_s(0, WU, sub)
            barrier()
            es[0].record(st_s)
            loop_s(WU, WU + K, sub)
            es[1].record(st_s)
            barrier()
            ms = max_over_ranks(es[0].elapsed_time(es[1]))
            strong[key] = tot * K / (ms * 1e-3)
            strong[key + "_ms_per_step"] = ms / K
        strong["unit"] = "frames/s"
        strong["note"] = ("%d sequences in total over %d GPUs; compare with the N = 1 line of the weak run, which is the same "
                          "%d sequences on one GPU" % (tot, world, tot))
        ctx_s.close()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    frames_total = total_streams * K
    stage_ms = stage / max(acc["n"], 1)                       # ingest+pyramid, LK, FAST (+mask), bucketing (per step, this rank)
    kp = kp_sum / (S * NS)
    n_trk = trk_sum / (S * NS)
    bk = bk_sum / (S * NS)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, peak_src = (peaks["hbm_gbs"], "MEASURED_PEAKS.json") if "hbm_gbs" in peaks else (6650.0, "fallback of B200_PROFILING.md")
    npyr = pyramid_pixels(w, h, levels)
    # algorithmic bytes per launch group (SURVEY.md section 8d x the streams one launch covers)
    alg = {
        "pyramid": S * npyr,                                  # read level 0 once + write levels >= 1
        "klt": S * (2 * npyr + 29 * n_trk),                   # both pyramids once + per-track in/out
        "fast": S * (w * h + 12 * kp),                        # image read + 12 B per keypoint (m = 0: the mask is evaluated from the tracked points, no mask plane exists)
        "bucket": S * 12 * (kp + bk),
    }
    names = ["pyramid", "klt", "fast", "bucket"]
    kernels = {n: {"ms": float(stage_ms[i]), "alg_bytes": float(alg[n]),
                   "achieved_gbs": float(alg[n] / (stage_ms[i] * 1e-3) / 1e9) if stage_ms[i] > 0 else None}
               for i, n in enumerate(names)}
    ncu = {}
    try:
        ncu = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except Exception:
        pass
    ncu_fresh = ncu.get("klt_source_sha") == source_sha()     # the capture is of THIS klt.cu
    klt = kernels["klt"]
    it_per_track = iters / max(S * n_trk * acc["n"], 1)
    # op model of SURVEY 8d: 18 lane-ops per setup pixel per level, 8 per iteration pixel
    lane_ops = (levels * wl["win"] ** 2 * 18 + it_per_track * wl["win"] ** 2 * 8) * n_trk * S
    sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
    lane_peak = 148 * 128 * sm_mhz * 1e6
    ach_ops = lane_ops / (klt["ms"] * 1e-3) if klt["ms"] > 0 else None
    kname = "klt_run_kernel<%d,%d,...>" % (wl["win"], wl["win"])
    roof = {"kernel": kname, "bound": "sm_issue", "achieved": ach_ops, "peak": lane_peak, "unit": "lane-op/s",
            "frac": (ach_ops or 0) / lane_peak,
            "traffic": ncu.get("klt_kernel") if ncu_fresh else None,
            "model": "SURVEY 8d: 18 lane-ops per set-up pixel per level + 8 per iteration pixel, measured iterations per track; "
                     "peak = 148 SMs x 128 lanes x the SM clock sampled during the timed region",
            "iterations_per_track": it_per_track, "lane_ops_per_launch": lane_ops, "launch_ms": klt["ms"],
            "hbm": {"achieved": klt["achieved_gbs"], "peak": hbm_peak, "unit": "GB/s", "frac": (klt["achieved_gbs"] or 0) / hbm_peak,
                    "peak_source": peak_src, "alg_bytes_per_launch": klt["alg_bytes"],
                    "note": "for the contract: LK's footprint (two pyramids) lives in L2; HBM is not its bound"}}
    if ncu_fresh and ncu.get("klt_warp_inst_per_launch") and S == 64 and wl["n_tracks"] == 2000 and klt["ms"] > 0:
        peak_issue = 148 * 4 * sm_mhz * 1e6
        ach = ncu["klt_warp_inst_per_launch"] / (klt["ms"] * 1e-3)
        roof["issue_slots"] = {"achieved": ach, "peak": peak_issue, "unit": "warp-inst/s", "frac": ach / peak_issue,
                               "warp_inst_per_launch": ncu["klt_warp_inst_per_launch"],
                               "warp_inst_per_track": ncu["klt_warp_inst_per_launch"] / (S * n_trk),
                               "source": "instructions: ncu smsp__inst_executed.sum of this klt.cu (profiles/ncu_traffic.json, "
                                         "source hash checked); time: live CUDA events"}
    line = {
        "metric": METRIC,
        "value": frames_total / (dev_ms * 1e-3), "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "strong" if args.total_streams else "weak",
        "vs_baseline": None, "dtype": DTYPE, "data": "synthetic", "config": workload_config(args, wl, world),
        "e2e": {"value": frames_total / (e2e_ms * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": S * w * h,
                "d2h_bytes_per_step": fe.d2h_bytes, "ms_per_step": e2e_ms / K},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roof,
        # FAST is limited by the integer ALU pipe (ncu pipe_alu 72 % of peak, profiles/r02_final_kernels.txt), not by HBM: its
        # GB/s figure is reported for completeness (bytes actually read: the image, no mask plane) and labelled accordingly
        "rooflines": {n: {"bound": "alu_pipe" if n == "fast" else "hbm", "achieved": kernels[n]["achieved_gbs"], "peak": hbm_peak,
                          "unit": "GB/s", "frac": (kernels[n]["achieved_gbs"] or 0) / hbm_peak, "ms": kernels[n]["ms"],
                          "alg_bytes_per_launch": kernels[n]["alg_bytes"],
                          **({"note": "frac is against HBM for comparison only; the limiting unit is the ALU pipe "
                                      "(ncu sm__inst_executed_pipe_alu 72 % of peak in the last capture)"} if n == "fast" else {})}
                      for n in ("pyramid", "fast", "bucket")},
        "kernels": kernels,
        "stats": {"keypoints_per_frame": kp, "bucketed_per_frame": bk, "tracks_per_frame": n_trk,
                  "status_ok_frac": ok_sum / max(trk_sum, 1), "streams_this_rank": S},
    }
    if strong:
        line["strong"] = strong
    line["klt_tracks_per_s"] = total_streams * n_trk * K / (dev_ms * 1e-3)
    # stage-only rates (KLT-only and FAST-only reported next to the whole front-end), this rank's launches
    line["stage_rates"] = {
        "klt_only_tracks_per_s": S * n_trk / (klt["ms"] * 1e-3) if klt["ms"] > 0 else None,
        "fast_only_frames_per_s": S / (kernels["fast"]["ms"] * 1e-3) if kernels["fast"]["ms"] > 0 else None,
        "pyramid_only_frames_per_s": S / (kernels["pyramid"]["ms"] * 1e-3) if kernels["pyramid"]["ms"] > 0 else None,
        "bucket_only_frames_per_s": S / (kernels["bucket"]["ms"] * 1e-3) if kernels["bucket"]["ms"] > 0 else None,
        "note": "per GPU; stage time = CUDA events around the stage's launches",
    }
    line["kernels"]["note"] = ("stage times are CUDA-event brackets over %d steps run one at a time right after the timed region "
                               "(one step outstanding: every bracket holds one stage alone; in the timed region the pyramid of "
                               "step t+1 runs beside FAST of step t); the FAST bracket includes the mask kernel" % NP_)
    line["notes"] = {"pipelining": "svo_gpu_step_submit / _wait, 2 steps outstanding; every step's results are waited for",
                     "untimed_steps_run": WU, "graphs": "the step's chain (LK .. counts) is replayed from captured CUDA graphs; "
                                                       "gpu_launches counts the kernels inside them",
                     "host_binding": binding}
    if h2d:
        line["h2d_probe"] = h2d
    if not args.no_extras and world == 1:
        # the reference's own track policy (svo_gpu_step_params.carry_tracks): LK survivors carry over, detection only on
        # key-frames (src/tracking.cpp:69-70, :168-177, :247-254); same streams, device-resident frames, 2 steps outstanding
        line["reference_loop"] = {"note": "carry_tracks mode, %d streams, frames resident in HBM; key-frame = the lane's first frame "
                                          "or a non-key-frame that ended below features_to_track; every step's results are waited for" % S}
        for ftt in (70, wl["n_tracks"]):
            try:
                fe2 = Frontend(ctx, w, h, S, bucket_capacity=cells, carry_tracks=1, features_to_track=ftt, **step_params(wl))
                n2, kf_sum, trk_sum2 = max(60, K // 3), 0, 0

                def sub2(i):
                    fe2.submit(dev[i % RING].data_ptr(), None, capi.MEM_DEVICE, pitch, h * pitch)

                def drain(i0, i1, count):
                    nonlocal kf_sum, trk_sum2
                    for i in range(i0, i1):
                        sub2(i)
                        if fe2.outstanding == 2:
                            fe2.wait()
                            if count:
                                kf_sum += int(fe2.keyframe.sum()); trk_sum2 += int(fe2.n_tracked.sum())
                    while fe2.outstanding:
                        fe2.wait()
                        if count:
                            kf_sum += int(fe2.keyframe.sum()); trk_sum2 += int(fe2.n_tracked.sum())

                fe2.reset()
                drain(0, 10, False)
                torch.cuda.synchronize()
                ev[4].record(stream)
                drain(10, 10 + n2, True)
                ev[5].record(stream)
                torch.cuda.synchronize()
                ms2 = ev[4].elapsed_time(ev[5]) / n2
                line["reference_loop"]["features_to_track_%d" % ftt] = {
                    "frames_per_s": S / (ms2 * 1e-3), "ms_per_step": ms2, "keyframe_fraction": kf_sum / (S * n2),
                    "tracks_per_frame": trk_sum2 / (S * n2), "steps": n2}
            except Exception as e:
                line["reference_loop"]["features_to_track_%d" % ftt] = {"error": repr(e)[:200]}
        fe.reset()
    if not args.no_extras:
        try:
            line["single_stream"] = run_single_stream(capi, Frontend, torch, local, wl, np.ascontiguousarray(frames[:, :1]), 200)
        except Exception as e:   # an extra must never cost the headline line
            line["single_stream"] = {"error": repr(e)[:200]}
        if world == 1:
            line["drop_in"] = run_drop_in(wl, np.ascontiguousarray(frames[:, 0]))
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_block(wl, 3, args.cpu_frames, args.cpu_single_frames)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
