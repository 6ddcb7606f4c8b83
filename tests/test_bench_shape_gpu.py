"""Parity of svo_gpu_step at the shape bench.py times: 64 lanes x 1241x376 x 2 000 tracks, mask on, bench seeds, the
persistent LK grid over a 128 000-entry work queue and the one-CTA-per-lane bucketing path.

Checked per lane and per frame, through the C ABI (submit / wait, two steps outstanding, host frames):
  FAST list (x, y, response, order)  bit-exact vs live OpenCV         (reference call site src/tracking.cpp:76-82)
  bucketing output                   bit-exact vs the reference's src/bucket.cpp compiled unmodified (:88, bucket.cpp:24-68)
  LK                                 <= 0.01 px and >= 99.9 % status agreement vs cv2.calcOpticalFlowPyrLK, every lane
                                     (src/tracking.cpp:160-165); bit-exact vs orc_klt on 8 lanes
and the literal C5 split: the same 64 sequences over two contexts of 32 lanes (sharding.sequences_of) give bit-identical
results to the one 64-lane context.
"""
import numpy as np
import pytest

from svo_b200 import capi, sharding, synth
from svo_b200.frontend import Frontend

pytestmark = pytest.mark.gpu

WL = synth.WORKLOADS["C1"]
W, H, S, T = WL["w"], WL["h"], 64, 4
KP_CAP = 24576
ORC_LANES = (0, 7, 13, 21, 34, 42, 55, 63)


def _params():
    return dict(fast_threshold=WL["fast_thr"], bucket_size=WL["bucket_size"], per_bucket=WL["per_bucket"],
                max_tracks=WL["n_tracks"], win_w=WL["win"], win_h=WL["win"], max_level=WL["max_level"], mask_half=10.0)


@pytest.fixture(scope="module")
def frames():
    """[T, S, H, W] uint8, stream s = bench.py's stream s of rank 0."""
    from concurrent.futures import ThreadPoolExecutor
    seeds = sharding.stream_seeds(0, S, WL["seed"] * 1000)
    with ThreadPoolExecutor(8) as ex:
        seqs = list(ex.map(lambda sd: synth.sequence(W, H, sd, T), seeds))
    return np.ascontiguousarray(np.stack(seqs, axis=1))


def _run(ctx, fr, batch):
    """Drives T frames with two steps outstanding; returns per frame a dict of copies of the host results."""
    import torch
    host = torch.from_numpy(fr).pin_memory()
    cells = (H // WL["bucket_size"] + 1) * (W // WL["bucket_size"] + 1) * WL["per_bucket"]
    fe = Frontend(ctx, W, H, batch, kp_capacity=KP_CAP, bucket_capacity=cells, **_params())
    res = []

    def collect():
        fe.wait()
        res.append({k: getattr(fe, k).copy() for k in
                    ("n_tracked", "n_keypoints", "n_bucketed", "next_xy", "status", "err", "kp_xy", "kp_response", "bucket_xy")})

    for t in range(T):
        nxt = host[t + 1].data_ptr() if t + 1 < T else None
        fe.submit(host[t].data_ptr(), nxt, capi.MEM_HOST, W, W * H)
        if fe.outstanding == 2:
            collect()
    while fe.outstanding:
        collect()
    return res, fe.pdict


@pytest.fixture(scope="module")
def gpu64(frames):
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=32768, max_batch=S) as ctx:
        res, pdict = _run(ctx, frames, S)
        assert ctx.last_klt_iterations > 0
    return res, pdict


def test_bench_shape_vs_opencv_and_reference_bucketing(oracle, frames, gpu64):
    from oracle import oracle as O
    if not (O.have_cv2() and O.have_ref_bucket()):
        pytest.skip("needs cv2 and oracle/_ref/libref_bucket.so")
    res, p = gpu64
    kw = dict(win=(p["win_w"], p["win_h"]), max_level=p["max_level"], max_count=p["max_count"], epsilon=p["epsilon"],
              criteria_type=p["criteria_type"], flags=p["flags"], min_eig_threshold=p["min_eig_threshold"])
    worst_px, worst_agree = 0.0, 1.0
    for s in range(S):
        tracks = np.zeros((0, 2), np.float32)
        for t in range(T):
            r, img = res[t], frames[t, s]
            # ---- LK from the previous frame on the tracks seeded by the previous step
            nt = int(r["n_tracked"][s])
            assert nt == len(tracks)
            if t > 0:
                assert nt == WL["n_tracks"]
                cxy, cst, _ = O.cv2_klt(frames[t - 1, s], img, tracks, **kw)
                gst = r["status"][s, :nt]
                agree = float(np.mean(gst == cst))
                both = (gst == 1) & (cst == 1)
                d = float(np.abs(r["next_xy"][s, :nt][both] - cxy[both]).max()) if both.any() else 0.0
                worst_px, worst_agree = max(worst_px, d), min(worst_agree, agree)
                assert agree >= 0.999, "lane %d frame %d: status agreement %.5f" % (s, t, agree)
                assert d <= 0.01, "lane %d frame %d: %.5f px" % (s, t, d)
                assert both.mean() > 0.9
                if s in ORC_LANES:
                    oxy, ost, oer = O.orc_klt(frames[t - 1, s], img, tracks, **kw)
                    assert np.array_equal(gst, ost)
                    assert np.array_equal(r["next_xy"][s, :nt], oxy)
                    assert np.array_equal(r["err"][s, :nt], oer)
            # ---- FAST masked around the points LK started from
            mask = O.cv2_build_mask(W, H, tracks, p["mask_half"])
            kxy, kresp = O.cv2_fast(img, p["fast_threshold"], True, mask)
            nk = int(r["n_keypoints"][s])
            assert nk == len(kxy) and nk <= KP_CAP
            assert np.array_equal(r["kp_xy"][s, :nk], kxy)
            assert np.array_equal(r["kp_response"][s, :nk], kresp)
            # ---- bucketing by the reference's own code
            bxy, _ = O.ref_bucket(kxy, np.zeros(len(kxy), np.int32), W, H, p["bucket_size"], p["per_bucket"])
            nb = int(r["n_bucketed"][s])
            assert nb == len(bxy)
            assert np.array_equal(r["bucket_xy"][s, :nb], bxy)
            assert nb >= WL["n_tracks"]          # the workload really carries 2 000 tracks per lane
            tracks = bxy[:p["max_tracks"]].copy()
    print("bench shape: worst LK deviation %.2e px, worst status agreement %.5f over %d lanes x %d frames" %
          (worst_px, worst_agree, S, T - 1))


def test_c5_split_over_two_contexts(frames, gpu64):
    """BASELINE config 5 as written: 64 sequences sharded over G = 2 contexts (what two ranks of bench.py
    --total-streams 64 own).  Every lane's results are bit-identical to the 64-lane run."""
    res64, _ = gpu64
    for rank in range(2):
        ids = sharding.sequences_of(rank, 2, S)
        assert len(ids) == S // 2
        with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=32768, max_batch=len(ids)) as ctx:
            res, _ = _run(ctx, np.ascontiguousarray(frames[:, ids]), len(ids))
        for t in range(T):
            for j, s in enumerate(ids):
                a, b = res[t], res64[t]
                for cnt, arrs in (("n_tracked", ("next_xy", "status", "err")), ("n_keypoints", ("kp_xy", "kp_response")),
                                  ("n_bucketed", ("bucket_xy",))):
                    n = int(a[cnt][j])
                    assert n == int(b[cnt][s])
                    if cnt == "n_tracked" and t == 0:
                        continue
                    for name in arrs:
                        assert np.array_equal(a[name][j, :n], b[name][s, :n]), (name, rank, s, t)


def test_graph_replay_at_bench_shape(frames, monkeypatch):
    """The step's chain replayed from captured CUDA graphs (default; six variants, recorded during steps 2 and 3, replayed from then on) against
    the same steps enqueued call by call (SVO_GPU_STEP_GRAPH=0), 64 lanes, 14 steps cycling the fixture's frames: every
    output of every lane and step bit-identical, and the same number of kernels counted."""
    import torch
    cells = (H // WL["bucket_size"] + 1) * (W // WL["bucket_size"] + 1) * WL["per_bucket"]
    host = torch.from_numpy(frames).pin_memory()
    runs = {}
    for graph in ("0", "1"):
        monkeypatch.setenv("SVO_GPU_STEP_GRAPH", graph)
        with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=32768, max_batch=S) as ctx:
            fe = Frontend(ctx, W, H, S, kp_capacity=KP_CAP, bucket_capacity=cells, **_params())
            res = []

            def collect():
                fe.wait()
                res.append({k: getattr(fe, k).copy() for k in ("n_tracked", "n_keypoints", "n_bucketed", "next_xy", "status", "err",
                                                             "kp_xy", "kp_response", "bucket_xy")})

            for t in range(14):
                fe.submit(host[t % T].data_ptr(), host[(t + 1) % T].data_ptr(), capi.MEM_HOST, W, W * H)
                if fe.outstanding == 2:
                    collect()
            while fe.outstanding:
                collect()
            runs[graph] = (res, ctx.launch_count)
    (a, la), (b, lb) = runs["0"], runs["1"]
    assert la == lb and len(a) == len(b) == 14
    for t in range(14):
        for cnt, arrs in (("n_tracked", ("next_xy", "status", "err")), ("n_keypoints", ("kp_xy", "kp_response")),
                          ("n_bucketed", ("bucket_xy",))):
            assert np.array_equal(a[t][cnt], b[t][cnt]), (t, cnt)
            if cnt == "n_tracked" and t == 0:
                continue
            for s in range(S):
                n = int(a[t][cnt][s])
                for name in arrs:
                    assert np.array_equal(a[t][name][s, :n], b[t][name][s, :n]), (t, s, name)
