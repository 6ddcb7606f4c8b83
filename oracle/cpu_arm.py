"""Timed CPU arm of bench.py (`cpu_baseline` and `--impl reference`).  TEST / MEASUREMENT INFRASTRUCTURE ONLY (see
oracle/oracle.py); svo_b200 never imports it.

What is timed is the reference's own CPU code path for one front-end frame (src/tracking.cpp:240-257 on a key-frame):
  cv::calcOpticalFlowPyrLK 21x21 (:160-165)            -> cv2.calcOpticalFlowPyrLK           (OpenCV itself)
  mask of +-10 px boxes (:76-79)                        -> orc_build_mask, a C loop           (bit-identical to the
                                                           cv::rectangle loop, tests/test_oracle_pin.py; 2 000 cv2.rectangle
                                                           calls from Python would charge ~2 ms of interpreter time per
                                                           frame to the CPU arm, which the reference's C++ loop does not pay)
  mDetector->detect(img, kps, mask) (:82)               -> cv2.FastFeatureDetector.detect      (OpenCV itself)
  cv::KeyPoint::convert (:86)                           -> cv2.KeyPoint_convert                (OpenCV itself, in C++)
  FeatureSet::bucketingFeatures (:88, src/bucket.cpp)   -> oracle/_ref/libref_bucket.so        (the reference's file, unmodified)

No per-keypoint Python runs inside the timed region.  Streams run in separate PROCESSES (spawn), one per host core, so no
interpreter lock is shared; each worker renders its own frames before the common start barrier.  Reported:
  t1          one sequence, cv2.setNumThreads(1)                  (frames/s, per-stage median ms over >= 50 frames)
  tall        one sequence, cv2.setNumThreads(all cores)          (the reference's actual configuration: one stream, OpenCV
                                                                   parallel_for_ inside LK)
  aggregate   one sequence per core, each cv2.setNumThreads(1)    (the best the host can do on many streams)
"""
import multiprocessing as mp
import os
import time

import numpy as np

STAGES = ("lk", "mask", "detect", "convert", "bucket")


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cv2_info():
    """cv2.getNumberOfCPUs() and the CPU-dispatch lines of cv2.getBuildInformation()."""
    import cv2
    keep = ("Baseline:", "requested:", "Dispatched code generation:", "Parallel framework:", "Intel IPP:")
    lines = [" ".join(l.split()) for l in cv2.getBuildInformation().splitlines() if any(k in l for k in keep)]
    return {"version": cv2.__version__, "getNumberOfCPUs": int(cv2.getNumberOfCPUs()), "build": lines}


class TimedFrontend:
    """One stream of the CPU front-end with per-stage wall-clock (same state machine as oracle.cpu_pipeline.CpuFrontend,
    whose answers tests/ compare with; this one keeps only what the reference's loop keeps)."""

    def __init__(self, p):
        import cv2
        from . import oracle as O
        self.cv2, self.O, self.p = cv2, O, dict(p)
        self.det = cv2.FastFeatureDetector_create(int(p["fast_threshold"]), bool(p["fast_nms"]))
        self.prev = None
        self.tracks = np.zeros((0, 2), np.float32)
        self.use_ref_bucket = O.have_ref_bucket()
        self.times = {k: [] for k in STAGES}
        self.total = []

    def step(self, img):
        cv2, O, p = self.cv2, self.O, self.p
        h, w = img.shape
        t0 = time.perf_counter()
        if self.prev is not None and len(self.tracks):
            cv2.calcOpticalFlowPyrLK(self.prev, img, self.tracks.reshape(-1, 1, 2), None, winSize=(p["win_w"], p["win_h"]),
                                     maxLevel=p["max_level"], criteria=(p["criteria_type"], p["max_count"], p["epsilon"]),
                                     flags=p["flags"], minEigThreshold=p["min_eig_threshold"])
        t1 = time.perf_counter()
        mask = None
        if p["mask_half"] > 0:
            mask = O.orc_build_mask(w, h, self.tracks if self.prev is not None else np.zeros((0, 2), np.float32), p["mask_half"])
        t2 = time.perf_counter()
        kps = self.det.detect(img, mask)
        t3 = time.perf_counter()
        xy = cv2.KeyPoint_convert(kps).reshape(-1, 2) if len(kps) else np.zeros((0, 2), np.float32)
        t4 = time.perf_counter()
        bucket = O.ref_bucket if self.use_ref_bucket else O.orc_bucket
        bxy, _ = bucket(xy, np.zeros(len(xy), np.int32), w, h, p["bucket_size"], p["per_bucket"])
        self.tracks = np.ascontiguousarray(bxy[:p["max_tracks"]])
        self.prev = img
        t5 = time.perf_counter()
        for k, dt in zip(STAGES, (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4)):
            self.times[k].append(dt)
        self.total.append(t5 - t0)


def _render(wl, seed, ring):
    from svo_b200 import synth
    times = synth.ring_times(ring)
    canvas = synth.make_canvas(wl["w"], wl["h"], seed)
    fr = {t: synth.render_frame(canvas, wl["w"], wl["h"], t) for t in sorted(set(times))}
    return [fr[t] for t in times]


def _worker(idx, wl, params, seed, ring, warmup, steps, cv_threads, barrier, queue):
    try:
        import cv2
        cv2.setNumThreads(cv_threads)
        frames = _render(wl, seed, ring)
        fe = TimedFrontend(params)
        for t in range(warmup):
            fe.step(frames[t % ring])
        for k in fe.times:
            fe.times[k].clear()
        fe.total.clear()
        barrier.wait(timeout=600)
        t0 = time.perf_counter()          # CLOCK_MONOTONIC: comparable across the processes of one host
        for t in range(warmup, warmup + steps):
            fe.step(frames[t % ring])
        t1 = time.perf_counter()
        queue.put((idx, t0, t1, {k: list(v) for k, v in fe.times.items()}, list(fe.total), "reference" if fe.use_ref_bucket else "port"))
    except Exception as e:   # the parent must not wait for ever
        try:
            barrier.abort()
        except Exception:
            pass
        queue.put((idx, None, None, repr(e), None, None))


def run(wl, params, seeds, warmup, steps, cv_threads, ring=8):
    """len(seeds) processes, one stream each, `warmup` untimed + `steps` timed frames per stream behind a common start.
    -> dict(frames_per_s, seconds, frames, processes, cv_threads, stage_ms_median, frame_ms_median, python_overhead_frac)."""
    from . import oracle as O
    O.build()                      # once, here: the workers only load what is already built
    ctx = mp.get_context("spawn")
    n = len(seeds)
    barrier, queue = ctx.Barrier(n), ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(i, wl, params, seeds[i], ring, warmup, steps, cv_threads, barrier, queue))
             for i in range(n)]
    for pr in procs:
        pr.start()
    res = [queue.get(timeout=900) for _ in range(n)]
    for pr in procs:
        pr.join(timeout=60)
    bad = sorted((r for r in res if r[1] is None), key=lambda r: "BrokenBarrier" in r[3])   # the root cause first
    if bad:
        raise RuntimeError("CPU arm worker failed: %s" % bad[0][3])
    t0, t1 = min(r[1] for r in res), max(r[2] for r in res)
    stage = {k: float(np.median(np.concatenate([np.asarray(r[3][k]) for r in res])) * 1e3) for k in STAGES}
    tot = np.concatenate([np.asarray(r[4]) for r in res])
    insum = sum(np.concatenate([np.asarray(r[3][k]) for r in res]).sum() for k in STAGES)
    # interpreter time between the five library calls of a frame (the calls themselves are C/C++ apart from argument
    # marshalling); loop overhead between frames is inside `seconds` but not inside `tot`
    wall = sum(r[2] - r[1] for r in res)
    return {"frames_per_s": n * steps / (t1 - t0), "seconds": t1 - t0, "frames": n * steps, "processes": n,
            "cv_threads": cv_threads, "stage_ms_median": stage, "frame_ms_median": float(np.median(tot) * 1e3),
            "python_overhead_frac": float(max(0.0, wall - insum) / wall), "kind": res[0][5]}


def full_report(wl, params, base_seed, warmup, steps, cores=None, single_frames=50):
    """The three figures of the module docstring.  `steps` / `warmup` size the aggregate run (the bench 'step' of the CPU
    arm = one frame of every stream); the single-sequence runs take >= single_frames frames each."""
    cores = cores or host_cores()
    agg = run(wl, params, [base_seed + i for i in range(cores)], warmup, steps, 1)
    t1 = run(wl, params, [base_seed], 3, max(single_frames, 1), 1)
    tall = run(wl, params, [base_seed], 3, max(single_frames, 1), cores)
    return {"aggregate": agg, "t1": t1, "tall": tall, "cores": cores, "cv2": cv2_info()}
