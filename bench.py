#!/usr/bin/env python
"""bench.py -- front-end frames/s of the svo per-frame feature front-end (FAST + bucketing + 4-level 21x21 KLT).

One "step" = one frame of every stream on this GPU through svo_gpu_step (libsvo_gpu.so, C ABI): ingest + pyramid,
LK from the previous frame on each stream's 2k tracks, FAST-9/16 + NMS on the new frame, grid bucketing, re-seeding of
the tracks.  Workload = BASELINE.json's metric configuration (1241x376 KITTI-shaped synthetic frames, 2k tracks per
stream) with `--streams` independent sequences per GPU (C5: 64).  Streams never interact, so N GPUs = N x streams
sequences with no collective ("scaling": "weak"); torch.distributed is used for the start/stop barriers only.

  value : frames/s with the frames already resident in HBM (device ring larger than L2), CUDA events on the context's stream
  e2e   : frames/s through the same C-ABI call with HOST frames (pinned), H2D of every frame and D2H of the step's
          results (tracked points, status, err, bucketed features, counts) inside the timed region
  --impl reference : the reference's own CPU path (OpenCV via cv2 + src/bucket.cpp compiled unmodified) on the host cores
"""
import argparse
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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--streams", type=int, default=64, help="independent sequences per GPU")
    ap.add_argument("--workload", default="C1", choices=sorted(synth.WORKLOADS), help="per-stream shape")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline sample")
    return ap.parse_args()


def make_frames(wl, streams, seed0, rank=0):
    """uint8 [RING, streams, h, w]: stream s = synthetic sequence with seed sharding.stream_seeds(rank, streams, seed0)[s]."""
    seeds = sharding.stream_seeds(rank, streams, seed0)
    w, h = wl["w"], wl["h"]
    times = synth.ring_times(RING)
    distinct = sorted(set(times))

    def one(s):
        canvas = synth.make_canvas(w, h, seeds[s])
        fr = {t: synth.render_frame(canvas, w, h, t) for t in distinct}
        return np.stack([fr[t] for t in times])

    with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as ex:
        seqs = list(ex.map(one, range(streams)))
    return np.ascontiguousarray(np.stack(seqs, axis=1))


def step_params(wl):
    return dict(fast_threshold=wl["fast_thr"], bucket_size=wl["bucket_size"], per_bucket=wl["per_bucket"],
                max_tracks=wl["n_tracks"], win_w=wl["win"], win_h=wl["win"], max_level=wl["max_level"])


def pyramid_pixels(w, h, levels):
    n = 0
    for _ in range(levels):
        n += w * h
        w, h = (w + 1) // 2, (h + 1) // 2
    return n


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "50"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self, t0=None, t1=None):
        """Summary of the samples taken between host times t0 and t1 (time.time()); all samples if none fall inside."""
        import datetime
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        rows = []
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                ts = datetime.datetime.strptime(c[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                rows.append((ts, float(c[1]), float(c[2]), float(c[3]), c[5:9]))
            except ValueError:
                continue
        os.unlink(self.f.name)
        inside = [r for r in rows if t0 is not None and t0 <= r[0] <= t1]
        use = inside or rows
        if use:
            reasons = set()
            for r in use:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            out.update(sm_mhz=float(np.median([r[1] for r in use])), sm_max_mhz=float(max(r[2] for r in use)),
                       power_w=float(np.median([r[3] for r in use])), reasons=sorted(reasons), samples=len(use),
                       window="timed region" if inside else "whole run (no sample fell inside the timed region)")
        return out


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_frontend_run(frames, wl, n_streams, n_steps, warmup, threads):
    """The reference's CPU path (cv2 + compiled src/bucket.cpp, else the C port) on `n_streams` streams for
    `warmup + n_steps` frames each, streams spread over `threads` host threads.  Returns (seconds of the timed steps,
    frames processed in them, kind)."""
    from oracle import oracle as O                     # bench.py's CPU arm is one of the places allowed to run oracle/
    from oracle.cpu_pipeline import CpuFrontend, default_params
    O.build()
    impl = "cv2" if O.have_cv2() else "orc"
    if impl == "cv2":
        import cv2
        cv2.setNumThreads(1 if n_streams >= threads else threads)   # parallelism across streams when there are enough
    p = default_params(wl["w"], wl["h"], wl["fast_thr"], wl["bucket_size"], wl["per_bucket"], wl["n_tracks"], wl["win"],
                       wl["max_level"])
    fes = [CpuFrontend(p, impl) for _ in range(n_streams)]
    pool = ThreadPoolExecutor(max_workers=max(1, min(threads, n_streams)))

    def run_step(t):
        list(pool.map(lambda s: fes[s].step(frames[t % RING, s]), range(n_streams)))

    for t in range(warmup):
        run_step(t)
    t0 = time.perf_counter()
    for t in range(warmup, warmup + n_steps):
        run_step(t)
    dt = time.perf_counter() - t0
    pool.shutdown()
    return dt, n_streams * n_steps, ("reference" if impl == "cv2" else "port")


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    wl = synth.WORKLOADS[args.workload]
    threads = os.cpu_count() or 1
    # bounded sample of the same workload: as many streams as host threads (each stream is sequential in time)
    n_streams = min(args.streams * args.gpus, max(threads, 1))
    frames = make_frames(wl, n_streams, wl["seed"] * 1000)
    dt, nframes, kind = cpu_frontend_run(frames, wl, n_streams, args.steps, args.warmup, threads)
    fps = nframes / dt
    line = {
        "impl": "reference", "metric": "front-end frames/s (FAST+bucketing+4-level 21x21 KLT, 2k tracks/stream)",
        "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8/int32 + f32 epilogue", "data": "synthetic",
        "config": workload_config(args, wl, note="CPU arm: %d streams per step on %d host threads" % (n_streams, threads)),
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind,
                         "sample": "%d streams x %d frames, cv2.setNumThreads(1), %d stream threads" %
                                   (n_streams, args.steps, min(threads, n_streams))},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(args, wl, note=None):
    c = {"workload": "C5-shaped: %d independent %dx%d synthetic sequences per GPU, %d tracks each (C1 per stream: FAST thr %d "
                     "+ NMS, bucketing %d px / %d per cell, KLT %dx%d, %d levels, 50 it / eps 0.001, min-eig err)" %
                     (args.streams, wl["w"], wl["h"], wl["n_tracks"], wl["fast_thr"], wl["bucket_size"], wl["per_bucket"],
                      wl["win"], wl["win"], wl["max_level"] + 1),
         "streams_per_gpu": args.streams, "width": wl["w"], "height": wl["h"], "tracks_per_stream": wl["n_tracks"],
         "parallelism": "independent sequences per GPU, no collective",
         "l2": "frame ring of %d x %d frames (%.0f MB) cycled, larger than the 126 MB L2" %
               (RING, args.streams, RING * args.streams * wl["h"] * 1280 / 1e6)}
    if note:
        c["note"] = note
    return c


# ------------------------------------------------------------------------------------------------ B200 arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    from svo_b200 import capi
    from svo_b200.frontend import Frontend

    rank, local, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the front-end has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's version banner (NCCL_DEBUG=VERSION on some boxes) goes to stderr
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "WARN"):
            os.environ.pop("NCCL_DEBUG")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = synth.WORKLOADS[args.workload]
    w, h, S, K, W = wl["w"], wl["h"], args.streams, args.steps, args.warmup
    levels = wl["max_level"] + 1
    pitch = (w + 127) // 128 * 128

    frames = make_frames(wl, S, wl["seed"] * 1000, rank)                  # [RING, S, h, w]
    dev = torch.zeros((RING, S, h, pitch), dtype=torch.uint8, device="cuda")
    dev[..., :w] = torch.from_numpy(frames).cuda()
    host = torch.from_numpy(frames).pin_memory()                         # compact rows, like a cv::Mat
    torch.cuda.synchronize()

    cells = (h // wl["bucket_size"] + 1) * (w // wl["bucket_size"] + 1) * wl["per_bucket"]
    ctx = capi.Context(device=local, max_w=w, max_h=h, max_levels=levels, max_points=max(32768, w * h // 16),
                       max_batch=S)
    fe = Frontend(ctx, w, h, S, bucket_capacity=cells, **step_params(wl))
    stream = torch.cuda.ExternalStream(ctx.stream)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(ms):
        return sharding.max_over_ranks(ms, world, device="cuda")

    # Both timed loops drive the step through svo_gpu_step_submit / svo_gpu_step_wait with two steps outstanding: step i+1 is
    # enqueued before the host waits for step i (frame i+1 never needs frame i's results on the HOST: the tracks are re-seeded
    # on the device), so the device does not idle while the host handles a frame.  Every step's results are waited for.
    def run_steps(i0, i1, submit, after_wait=None):
        for i in range(i0, i1):
            submit(i)
            if fe.outstanding == 2:
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

    fe.reset()
    run_steps(0, W, submit_host)
    barrier()
    ev[0].record(stream)
    run_steps(W, W + K, submit_host)
    ev[1].record(stream)
    barrier()
    e2e_ms = max_over_ranks(ev[0].elapsed_time(ev[1]))
    # statistics of the workload (outside any timed region)
    kp_sum = bk_sum = trk_sum = ok_sum = 0
    NS = 8
    for i in range(W + K, W + K + NS):
        fe.step_host_pipelined(hptr(i), hptr(i + 1), w, w * h)
        kp_sum += int(fe.n_keypoints.sum()); bk_sum += int(fe.n_bucketed.sum()); trk_sum += int(fe.n_tracked.sum())
        ok_sum += int(sum(int(fe.status[b, :fe.n_tracked[b]].sum()) for b in range(S)))

    # ---- device resident: frames already in HBM, nothing returned to the host
    sampler = ClockSampler(local) if rank == 0 else None   # nvidia-smi needs a moment to start: launched before the warm-up
    def submit_dev(i):
        fe.submit(dev[i % RING].data_ptr(), None, capi.MEM_DEVICE, pitch, h * pitch, with_outputs=False)

    fe.reset()
    run_steps(0, W, submit_dev)
    ctx.step_profile(True)
    launches0 = ctx.launch_count
    barrier()
    t_start = time.time()
    ev[2].record(stream)
    acc = {"stage": np.zeros(4), "iters": 0}

    def after_wait():
        acc["stage"] += np.array(ctx.step_profile(True))
        acc["iters"] += ctx.last_klt_iterations

    run_steps(W, W + K, submit_dev, after_wait)
    stage, iters = acc["stage"], acc["iters"]
    ev[3].record(stream)
    barrier()
    clocks = sampler.stop(t_start, time.time()) if sampler else None
    dev_ms = max_over_ranks(ev[2].elapsed_time(ev[3]))
    launches = ctx.launch_count - launches0
    ctx.step_profile(False)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    frames_total = world * S * K
    stage_ms = stage / K                                      # ingest+pyramid, LK, FAST, bucketing (per step, this rank)
    kp = kp_sum / (S * NS)
    n_trk = trk_sum / (S * NS)
    bk = bk_sum / (S * NS)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    npyr = pyramid_pixels(w, h, levels)
    # algorithmic bytes per launch group (SURVEY.md section 8d x the streams one launch covers)
    alg = {
        "pyramid": S * npyr,                                  # read level 0 once + write levels >= 1
        "klt": S * (2 * npyr + 29 * n_trk),                   # both pyramids once + per-track in/out
        "fast": S * (w * h + 12 * kp),
        "bucket": S * 12 * (kp + bk),
    }
    names = ["pyramid", "klt", "fast", "bucket"]
    kernels = {n: {"ms": float(stage_ms[i]), "alg_bytes": float(alg[n]),
                   "achieved_gbs": float(alg[n] / (stage_ms[i] * 1e-3) / 1e9) if stage_ms[i] > 0 else None}
               for i, n in enumerate(names)}
    traffic = ncu = None
    try:
        ncu = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        traffic = ncu.get("klt_kernel")
    except Exception:
        pass
    klt = kernels["klt"]
    it_per_track = iters / max(S * n_trk * K, 1)
    # op model of SURVEY 8d: 18 lane-ops per setup pixel per level, 8 per iteration pixel
    lane_ops = (levels * wl["win"] ** 2 * 18 + it_per_track * wl["win"] ** 2 * 8) * n_trk * S
    sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
    line = {
        "metric": "front-end frames/s (FAST+bucketing+4-level 21x21 KLT, 2k tracks/stream)",
        "value": frames_total / (dev_ms * 1e-3), "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8/int32 + f32 epilogue", "data": "synthetic", "config": workload_config(args, wl),
        "e2e": {"value": frames_total / (e2e_ms * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": S * w * h,
                "d2h_bytes_per_step": fe.d2h_bytes, "ms_per_step": e2e_ms / K},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"kernel": "klt_kernel<21,21,8>", "bound": "hbm", "achieved": klt["achieved_gbs"], "peak": hbm_peak,
                     "unit": "GB/s", "frac": (klt["achieved_gbs"] or 0) / hbm_peak, "traffic": traffic,
                     "peak_source": peak_src,
                     "note": "LK is not HBM-bound: its footprint (two pyramids) lives in L2; its bound is the SM issue "
                             "rate, see roofline.issue and klt_issue_model"},
        "klt_issue_model": {"iterations_per_track": it_per_track, "lane_ops_per_launch": lane_ops,
                            "achieved_lane_ops_per_s": lane_ops / (klt["ms"] * 1e-3) if klt["ms"] > 0 else None,
                            "peak_lane_ops_per_s": 148 * 128 * sm_mhz * 1e6,
                            "frac": (lane_ops / (klt["ms"] * 1e-3)) / (148 * 128 * sm_mhz * 1e6) if klt["ms"] > 0 else None},
        "kernels": kernels,
        "stats": {"keypoints_per_frame": kp, "bucketed_per_frame": bk, "tracks_per_frame": n_trk,
                  "status_ok_frac": ok_sum / max(trk_sum, 1)},
    }
    line["klt_tracks_per_s"] = world * S * n_trk * K / (dev_ms * 1e-3)
    # stage-only rates (KLT-only and FAST-only reported next to the whole front-end), this rank's launches
    line["stage_rates"] = {
        "klt_only_tracks_per_s": S * n_trk / (klt["ms"] * 1e-3) if klt["ms"] > 0 else None,
        "fast_only_frames_per_s": S / (kernels["fast"]["ms"] * 1e-3) if kernels["fast"]["ms"] > 0 else None,
        "pyramid_only_frames_per_s": S / (kernels["pyramid"]["ms"] * 1e-3) if kernels["pyramid"]["ms"] > 0 else None,
        "bucket_only_frames_per_s": S / (kernels["bucket"]["ms"] * 1e-3) if kernels["bucket"]["ms"] > 0 else None,
        "note": "per GPU; stage time = CUDA events around the stage's launches inside the timed region",
    }
    line["kernels"]["note"] = ("stage times are live CUDA-event brackets inside the timed region; the pyramid of step t+1 is "
                               "built on a second stream beside FAST and bucketing of step t (SVO_GPU_STEP_PYR_ASYNC=1), so the "
                               "pyramid and FAST brackets overlap each other and their sum exceeds ms_per_step; LK runs alone")
    line["config"]["pipelining"] = "svo_gpu_step_submit / _wait, 2 steps outstanding; every step's results are waited for"
    # the bound LK actually runs against: warp instructions issued per second.  Instructions per launch come from the ncu
    # capture of this same workload (profiles/ncu_traffic.json, 64 x 2000 tracks), the launch time is measured here.
    if ncu and ncu.get("klt_warp_inst_per_launch") and S == 64 and wl["n_tracks"] == 2000 and klt["ms"] > 0:
        peak_issue = 148 * 4 * sm_mhz * 1e6
        ach = ncu["klt_warp_inst_per_launch"] / (klt["ms"] * 1e-3)
        line["roofline"]["issue"] = {"bound": "sm issue slots", "achieved": ach, "peak": peak_issue, "unit": "warp-inst/s",
                                     "frac": ach / peak_issue, "warp_inst_per_launch": ncu["klt_warp_inst_per_launch"],
                                     "source": "instructions: ncu smsp__inst_executed.sum (profiles/); time: live CUDA events"}
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        ns = min(S, threads)
        nsteps = max(4, int(args.cpu_seconds / 0.06 / ns))
        dt, nfr, kind = cpu_frontend_run(frames, wl, ns, nsteps, 2, threads)
        line["cpu_baseline"] = {"value": nfr / dt, "unit": "frames/s", "cores": threads, "kind": kind,
                                "sample": "%d streams x %d frames of the same workload, one stream per host thread" % (ns, nsteps)}
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
