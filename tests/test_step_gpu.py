"""GPU parity of the batched per-frame step (svo_gpu_step) against the CPU pipeline built from the oracle's functions:
LK from the previous frame, FAST (+ tracked-points mask), bucketing, track re-seeding -- bit-exact, per lane."""
import numpy as np
import pytest

from svo_b200 import capi, synth
from svo_b200.frontend import Frontend

pytestmark = pytest.mark.gpu

W, H = 320, 240


def _seqs(n_frames=4):
    return np.stack([synth.sequence(W, H, seed, n_frames) for seed in (5, 6, 7)])   # [lane, t, h, w]


def _check_lane(fe, lane, ref, first):
    nt = int(fe.n_tracked[lane])
    assert nt == len(ref["next_xy"])
    if not first:
        assert nt > 100
        assert np.array_equal(fe.status[lane, :nt], ref["status"])
        assert np.array_equal(fe.next_xy[lane, :nt], ref["next_xy"])
        assert np.array_equal(fe.err[lane, :nt], ref["err"])
    nk = int(fe.n_keypoints[lane])
    assert nk == len(ref["kp_xy"])
    assert np.array_equal(fe.kp_xy[lane, :nk], ref["kp_xy"])
    assert np.array_equal(fe.kp_response[lane, :nk], ref["kp_response"])
    nb = int(fe.n_bucketed[lane])
    assert nb == len(ref["bucket_xy"])
    assert np.array_equal(fe.bucket_xy[lane, :nb], ref["bucket_xy"])


@pytest.mark.parametrize("mask_half", [0.0, 10.0])
def test_step_matches_cpu_pipeline(oracle, mask_half):
    from oracle.cpu_pipeline import CpuFrontend
    seqs = _seqs()
    B = seqs.shape[0]
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=B) as ctx:
        fe = Frontend(ctx, W, H, B, kp_capacity=8192, bucket_capacity=2048, max_tracks=400, mask_half=mask_half)
        cpus = [CpuFrontend(fe.pdict, impl="orc") for _ in range(B)]
        for t in range(seqs.shape[1]):
            fe.step_host(np.ascontiguousarray(seqs[:, t]))
            for b in range(B):
                _check_lane(fe, b, cpus[b].step(seqs[b, t]), first=(t == 0))
        assert ctx.last_klt_iterations > 0


@pytest.mark.parametrize("overlap", ["1", "0"])
def test_step_detection_beside_tracking(oracle, monkeypatch, overlap):
    """SVO_GPU_STEP_OVERLAP (read by svo_gpu_create): 1 (the default, as in every other test) = FAST -> bucketing on a second
    stream beside the LK kernel, 0 = back to back on one stream.  Same answers.  The last steps are profiled: a profiled step
    always runs its stages back to back, so that each bracket holds one stage."""
    from oracle.cpu_pipeline import CpuFrontend
    monkeypatch.setenv("SVO_GPU_STEP_OVERLAP", overlap)
    seqs = _seqs(6)
    B = seqs.shape[0]
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=B) as ctx:
        fe = Frontend(ctx, W, H, B, kp_capacity=8192, bucket_capacity=2048, max_tracks=400, mask_half=10.0)
        cpus = [CpuFrontend(fe.pdict, impl="orc") for _ in range(B)]
        for t in range(seqs.shape[1]):
            if t == 4:
                ctx.step_profile(True)
            fe.step_host(np.ascontiguousarray(seqs[:, t]))
            for b in range(B):
                _check_lane(fe, b, cpus[b].step(seqs[b, t]), first=(t == 0))
        ms = ctx.step_profile(False)
        assert all(m > 0 for m in ms) and ctx.last_klt_iterations > 0


def test_step_pyramid_on_main_stream(oracle, monkeypatch):
    """SVO_GPU_STEP_PYR_ASYNC=0: the pyramid is built on the main stream instead of the high-priority one (the default, which
    every other test of this file exercises).  Same answers, with two steps outstanding."""
    import torch
    from oracle.cpu_pipeline import CpuFrontend
    monkeypatch.setenv("SVO_GPU_STEP_PYR_ASYNC", "0")
    seqs = _seqs(5)
    B, T = seqs.shape[0], seqs.shape[1]
    frames = torch.from_numpy(np.ascontiguousarray(seqs.transpose(1, 0, 2, 3))).pin_memory()
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=B) as ctx:
        fe = Frontend(ctx, W, H, B, kp_capacity=8192, bucket_capacity=2048, max_tracks=400, mask_half=10.0)
        cpus = [CpuFrontend(fe.pdict, impl="orc") for _ in range(B)]
        fe.submit(frames[0].data_ptr(), None, capi.MEM_HOST, W, W * H)
        for t in range(1, T + 1):
            if t < T:
                fe.submit(frames[t].data_ptr(), None, capi.MEM_HOST, W, W * H)
            fe.wait()
            for b in range(B):
                _check_lane(fe, b, cpus[b].step(seqs[b, t - 1]), first=(t == 1))


def test_step_device_memory_and_reset(oracle):
    import torch
    from oracle.cpu_pipeline import CpuFrontend
    seqs = _seqs(3)
    B = seqs.shape[0]
    pitch = 384                                          # padded device rows
    dev = torch.zeros((seqs.shape[1], B, H, pitch), dtype=torch.uint8, device="cuda")
    dev[:, :, :, :W] = torch.from_numpy(np.ascontiguousarray(seqs.transpose(1, 0, 2, 3))).cuda()
    torch.cuda.synchronize()
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=B) as ctx:
        fe = Frontend(ctx, W, H, B, kp_capacity=8192, bucket_capacity=2048, max_tracks=400)
        for rep in range(2):                             # second pass after a reset gives the same answers
            fe.reset()
            cpus = [CpuFrontend(fe.pdict, impl="orc") for _ in range(B)]
            for t in range(seqs.shape[1]):
                fe.step_device(dev[t].data_ptr(), pitch, H * pitch, with_outputs=True)
                for b in range(B):
                    _check_lane(fe, b, cpus[b].step(seqs[b, t]), first=(t == 0))


def test_step_pipelined_ingest(oracle):
    """svo_gpu_step_pipelined: the next step's host frames are copied while this step computes; same answers, including
    when a prefetch is not used (different pointer) and after a reset."""
    import torch
    from oracle.cpu_pipeline import CpuFrontend
    seqs = _seqs(5)
    B, T = seqs.shape[0], seqs.shape[1]
    host = torch.from_numpy(np.ascontiguousarray(seqs.transpose(1, 0, 2, 3))).pin_memory()   # [t, lane, h, w]
    other = host[2].clone().pin_memory()
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=B) as ctx:
        fe = Frontend(ctx, W, H, B, kp_capacity=8192, bucket_capacity=2048, max_tracks=400)
        for rep in range(2):
            fe.reset()
            cpus = [CpuFrontend(fe.pdict, impl="orc") for _ in range(B)]
            for t in range(T):
                nxt = host[t + 1].data_ptr() if t + 1 < T else None
                cur = host[t].data_ptr()
                if t == 1:
                    nxt = host[4].data_ptr()          # a prefetch that the next call does not use
                if t == 2:
                    cur = other.data_ptr()            # same pixels as host[2], different pointer: copied normally
                fe.step_host_pipelined(cur, nxt, W, W * H)
                for b in range(B):
                    _check_lane(fe, b, cpus[b].step(seqs[b, t]), first=(t == 0))


@pytest.mark.parametrize("graph", ["1", "0"])
@pytest.mark.parametrize("mem", ["host", "device"])
def test_step_submit_wait(oracle, mem, graph, monkeypatch):
    """svo_gpu_step_submit / svo_gpu_step_wait with two steps outstanding (submit(t + 1) before wait(t)): each step's
    results, read after ITS wait, equal the synchronous CPU pipeline; host frames are prefetched one step ahead.
    SVO_GPU_STEP_GRAPH=1: from the third step on the chain LK .. counts is a captured CUDA graph, replayed per step (the next
    step's pyramid build waits for none of it: it writes the free slot of the three-slot ring); 0: every step enqueued call
    by call."""
    import torch
    from oracle.cpu_pipeline import CpuFrontend
    monkeypatch.setenv("SVO_GPU_STEP_GRAPH", graph)
    seqs = _seqs(9)
    B, T = seqs.shape[0], seqs.shape[1]
    frames = torch.from_numpy(np.ascontiguousarray(seqs.transpose(1, 0, 2, 3)))           # [t, lane, h, w]
    frames = frames.pin_memory() if mem == "host" else frames.cuda()
    kind = capi.MEM_HOST if mem == "host" else capi.MEM_DEVICE
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=B) as ctx:
        fe = Frontend(ctx, W, H, B, kp_capacity=8192, bucket_capacity=2048, max_tracks=400, mask_half=10.0)
        cpus = [CpuFrontend(fe.pdict, impl="orc") for _ in range(B)]

        def submit(t):
            nxt = frames[t + 1].data_ptr() if (mem == "host" and t + 1 < T) else None
            return fe.submit(frames[t].data_ptr(), nxt, kind, W, W * H)

        def check(t):
            fe.wait()
            for b in range(B):
                _check_lane(fe, b, cpus[b].step(seqs[b, t]), first=(t == 0))

        submit(0)
        for t in range(1, T):
            submit(t)
            assert fe.outstanding == 2
            check(t - 1)
        check(T - 1)
        assert fe.outstanding == 0 and ctx.last_klt_iterations > 0
        # launches are counted the same way replayed or enqueued
        assert ctx.launch_count > 0


def test_step_submit_wait_errors():
    seqs = _seqs(1)
    img = np.ascontiguousarray(seqs[:, 0])
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=3) as ctx:
        fe = Frontend(ctx, W, H, 3, kp_capacity=0, bucket_capacity=2048, max_tracks=400)
        t0 = fe.submit(img.ctypes.data, None, capi.MEM_HOST, W, W * H)
        t1 = fe.submit(img.ctypes.data, None, capi.MEM_HOST, W, W * H)
        with pytest.raises(capi.SvoGpuError) as e:                       # a third step while two are outstanding
            ctx.step_submit(img.ctypes.data, None, capi.MEM_HOST, W, W * H, 3, fe.params, None)
        assert e.value.code == capi.E_STATE
        ctx.step_wait(t0)
        with pytest.raises(capi.SvoGpuError) as e:                       # waited for twice
            ctx.step_wait(t0)
        assert e.value.code == capi.E_STATE
        with pytest.raises(capi.SvoGpuError):                            # never issued
            ctx.step_wait(t1 + 5)
        ctx.step_wait(t1)
        fe.reset()                                                       # a reset drops what is outstanding
        t2 = fe.submit(img.ctypes.data, None, capi.MEM_HOST, W, W * H)
        fe.reset()
        with pytest.raises(capi.SvoGpuError):
            ctx.step_wait(t2)
        fe.step_host(img)                                                # and the synchronous call still works
        assert int(fe.n_keypoints.sum()) > 0


def test_two_contexts_interleaved(oracle):
    """Two contexts on one device (two host threads' worth of state) driven alternately, with page-locked frames from
    svo_gpu_host_alloc: neither disturbs the other (streams, events, work counters and buffers are per context)."""
    import ctypes as C
    from oracle.cpu_pipeline import CpuFrontend
    seqs = _seqs(4)
    B, T = seqs.shape[0], seqs.shape[1]
    lib = capi.load()
    p = C.c_void_p()
    nbytes = B * W * H
    assert lib.svo_gpu_host_alloc(C.byref(p), T * nbytes) == 0 and p.value
    try:
        host = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(T, B, H, W))
        host[...] = seqs.transpose(1, 0, 2, 3)
        with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=B) as c1, \
                capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=B) as c2:
            f1 = Frontend(c1, W, H, B, kp_capacity=8192, bucket_capacity=2048, max_tracks=400, mask_half=10.0)
            f2 = Frontend(c2, W, H, B, kp_capacity=8192, bucket_capacity=2048, max_tracks=300, mask_half=0.0)
            cp1 = [CpuFrontend(f1.pdict, impl="orc") for _ in range(B)]
            cp2 = [CpuFrontend(f2.pdict, impl="orc") for _ in range(B)]
            for t in range(T):
                f1.submit(p.value + t * nbytes, None, capi.MEM_HOST, W, W * H)
                f2.submit(p.value + (T - 1 - t) * nbytes, None, capi.MEM_HOST, W, W * H)     # the sequences backwards
                f2.wait()
                f1.wait()
                for b in range(B):
                    _check_lane(f1, b, cp1[b].step(seqs[b, t]), first=(t == 0))
                    _check_lane(f2, b, cp2[b].step(seqs[b, T - 1 - t]), first=(t == 0))
    finally:
        assert lib.svo_gpu_host_free(p) == 0
    assert lib.svo_gpu_host_alloc(None, 16) != 0 and lib.svo_gpu_host_free(None) == 0


def test_step_capacity_error():
    seqs = _seqs(1)
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=3) as ctx:
        fe = Frontend(ctx, W, H, 3, kp_capacity=0, bucket_capacity=16, max_tracks=400)
        with pytest.raises(capi.SvoGpuError) as e:
            fe.step_host(np.ascontiguousarray(seqs[:, 0]))
        assert e.value.code == capi.E_CAPACITY


def test_step_with_pageable_outputs(oracle, monkeypatch):
    """Result arrays in ordinary (pageable) host memory: device-to-host copies into them cannot be captured, so the step keeps
    enqueueing call by call instead of replaying a graph -- same answers, two steps outstanding."""
    from oracle.cpu_pipeline import CpuFrontend
    from svo_b200 import frontend
    monkeypatch.setattr(frontend, "_pinned", lambda shape, dtype: (np.empty(shape, dtype), None))
    seqs = _seqs(7)
    B, T = seqs.shape[0], seqs.shape[1]
    frames = np.ascontiguousarray(seqs.transpose(1, 0, 2, 3))
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=B) as ctx:
        fe = Frontend(ctx, W, H, B, kp_capacity=8192, bucket_capacity=2048, max_tracks=400, mask_half=10.0)
        cpus = [CpuFrontend(fe.pdict, impl="orc") for _ in range(B)]
        done = 0
        for t in range(T):
            fe.submit(frames[t].ctypes.data, None, capi.MEM_HOST, W, W * H)
            if fe.outstanding == 2:
                fe.wait()
                for b in range(B):
                    _check_lane(fe, b, cpus[b].step(seqs[b, done]), first=(done == 0))
                done += 1
        while fe.outstanding:
            fe.wait()
            for b in range(B):
                _check_lane(fe, b, cpus[b].step(seqs[b, done]), first=(done == 0))
            done += 1
        assert done == T


def test_step_graph_survives_bucket_table_regrowth(oracle, monkeypatch):
    """A captured step graph holds the bucket tables' addresses by value.  svo_gpu_bucket with a finer grid on the same
    context frees and re-allocates those tables; the next step must record its graph again (allocation epoch in the graph
    key) instead of replaying one that writes freed memory.  SVO_GPU_BUCKET_MODE=1 keeps the step on the three
    global-memory kernels, the only path that uses the tables."""
    from oracle.cpu_pipeline import CpuFrontend
    monkeypatch.setenv("SVO_GPU_BUCKET_MODE", "1")
    monkeypatch.setenv("SVO_GPU_STEP_GRAPH", "1")
    seqs = _seqs(10)
    B = seqs.shape[0]
    rng = np.random.default_rng(12)
    pts = rng.uniform([0, 0], [W, H], (3000, 2)).astype(np.float32)
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=B) as ctx:
        fe = Frontend(ctx, W, H, B, kp_capacity=8192, bucket_capacity=2048, max_tracks=400, mask_half=10.0)
        cpus = [CpuFrontend(fe.pdict, impl="orc") for _ in range(B)]
        for t in range(seqs.shape[1]):
            if t == 6:                                   # graphs of all slots are recorded by now; grow the tables under them
                gx, ga = ctx.bucket(pts, np.zeros(len(pts), np.int32), W, H, 3, 6)
                ox, oa = oracle.orc_bucket(pts, np.zeros(len(pts), np.int32), W, H, 3, 6)
                assert np.array_equal(gx, ox) and np.array_equal(ga, oa)
            fe.step_host(np.ascontiguousarray(seqs[:, t]))
            for b in range(B):
                _check_lane(fe, b, cpus[b].step(seqs[b, t]), first=(t == 0))
