"""The host C++ wrappers with the reference's call signatures (svo_b200/host: GpuFastDetector::detect,
svo::calcOpticalFlowPyrLK, FeatureSet::bucketingFeatures), driven by svo_b200/lib/frontend_selftest the way
Tracking::extractFeatures / findLeftFeaturesInRight / trackFrames drive OpenCV (src/tracking.cpp:74-118, :154-179).
CPU: the wrappers build, and without a device they throw instead of falling back.  GPU: results equal the oracle's."""
import os
import subprocess

import numpy as np
import pytest

from svo_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "svo_b200", "lib", "frontend_selftest")


@pytest.fixture(scope="module")
def exe():
    if not os.path.exists(EXE):
        import __graft_entry__ as g
        g.build()
    return EXE


def test_wrappers_throw_without_gpu(exe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = subprocess.run([exe, "--expect-no-gpu"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "threw as expected" in r.stdout


def test_reference_signatures_declared():
    hdr = open(os.path.join(ROOT, "svo_b200", "host", "svo_frontend.hpp")).read()
    bk = open(os.path.join(ROOT, "svo_b200", "host", "bucket.h")).read()
    for needle in ("class GpuFastDetector : public cv::Feature2D", "void calcOpticalFlowPyrLK(", "cv::Size winSize = cv::Size(21, 21)",
                   "int maxLevel = 3", "int flags = 0, double minEigThreshold = 1e-4"):
        assert needle in hdr, needle
    for needle in ("struct FeatureSet", "std::vector<cv::Point2f> points;", "std::vector<int> ages;",
                   "void appendNewFeatures(const std::vector<cv::Point2f>& features)",
                   "void bucketingFeatures(const cv::Mat& image, int bucket_size, int features_per_bucket);"):
        assert needle in bk, needle


@pytest.mark.gpu
def test_sequence_batch_matches_cpu_pipeline(exe, oracle, tmp_path):
    """svo::SequenceBatch (C++ driver of svo_gpu_step_submit / _wait, two frame sets outstanding) against the CPU pipeline
    built from the oracle, per sequence and per frame, bit-exact."""
    from oracle.cpu_pipeline import CpuFrontend
    from svo_b200.frontend import make_params
    w, h, S, T, max_tracks = 320, 240, 3, 5, 400
    seqs = np.stack([synth.sequence(w, h, seed, T) for seed in (5, 6, 7)])      # [stream, t, h, w]
    fin, fout = tmp_path / "bin.bin", tmp_path / "bout.bin"
    with open(fin, "wb") as f:
        np.array([w, h, S, T, max_tracks], np.int32).tofile(f)
        np.ascontiguousarray(seqs.transpose(1, 0, 2, 3)).tofile(f)
    r = subprocess.run([exe, "--batch", str(fin), str(fout)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    buf = open(fout, "rb").read()
    off = 0

    def take(dtype, n):
        nonlocal off
        out = np.frombuffer(buf, dtype, n, off)
        off += out.nbytes
        return out

    _, pdict = make_params(w, h, max_tracks=max_tracks, mask_half=10.0)          # svo::FrontendParams defaults
    cpus = [CpuFrontend(pdict, impl="orc") for _ in range(S)]
    for t in range(T):
        for s in range(S):
            ref = cpus[s].step(seqs[s, t])
            nt = int(take(np.int32, 1)[0])
            xy, st, er = take(np.float32, 2 * nt).reshape(nt, 2), take(np.uint8, nt), take(np.float32, nt)
            nb = int(take(np.int32, 1)[0])
            feats = take(np.float32, 2 * nb).reshape(nb, 2)
            nk = int(take(np.int32, 1)[0])
            assert nt == len(ref["next_xy"]) and (t == 0 or nt > 100), (t, s)
            assert np.array_equal(xy, ref["next_xy"]) and np.array_equal(st, ref["status"]) and np.array_equal(er, ref["err"]), (t, s)
            assert nk == len(ref["kp_xy"]) and np.array_equal(feats, ref["bucket_xy"]), (t, s)
    assert int(take(np.int32, 1)[0]) == 1 and off == len(buf)


@pytest.mark.gpu
@pytest.mark.parametrize("cache", ["1", "0"])
def test_wrappers_match_oracle(exe, oracle, tmp_path, cache):
    """cache = the wrappers' content-addressed frame cache (default on; SVO_FRAME_CACHE=0 uploads and rebuilds per call, as
    cv::calcOpticalFlowPyrLK itself does): same answers either way."""
    w, h, bs, k = 320, 240, 32, 5
    canvas = synth.make_canvas(w, h, 5)
    a, b = synth.render_frame(canvas, w, h, 0), synth.render_frame(canvas, w, h, 1)
    prev = np.random.default_rng(1).uniform([0, 0], [w, h], (25, 2)).astype(np.float32)
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    with open(fin, "wb") as f:
        np.array([w, h, len(prev), bs, k], np.int32).tofile(f)
        a.tofile(f)
        b.tofile(f)
        prev.tofile(f)
    r = subprocess.run([exe, str(fin), str(fout)], capture_output=True, text=True, env=dict(os.environ, SVO_FRAME_CACHE=cache))
    assert r.returncode == 0, r.stderr
    buf = open(fout, "rb").read()
    off = 0

    def take(dtype, n):
        nonlocal off
        out = np.frombuffer(buf, dtype, n, off)
        off += out.nbytes
        return out

    K = int(take(np.int32, 1)[0])
    kp = take(np.float32, 3 * K).reshape(K, 3)
    mask = oracle.orc_build_mask(w, h, prev, 10.0)
    oxy, oresp = oracle.orc_fast(a, 20, True, mask=mask)
    assert K == len(oxy) and K > 100
    assert np.array_equal(kp[:, :2], oxy) and np.array_equal(kp[:, 2], oresp)
    assert int(take(np.int32, 1)[0]) == 1                    # device-built mask gave the same keypoints
    nb = int(take(np.int32, 1)[0])
    bxy = take(np.float32, 2 * nb).reshape(nb, 2)
    bag = take(np.int32, nb)
    ox, oa = oracle.orc_bucket(oxy, np.zeros(K, np.int32), w, h, bs, k)
    assert np.array_equal(bxy, ox) and np.array_equal(bag, oa)
    for mode in (dict(win=(11, 11), max_level=3, max_count=30, epsilon=0.001, flags=0),
                 dict(win=(21, 21), max_level=3, max_count=50, epsilon=0.001, flags=8)):
        n = int(take(np.int32, 1)[0])
        assert n == nb
        nxy = take(np.float32, 2 * n).reshape(n, 2)
        st = take(np.uint8, n)
        er = take(np.float32, n)
        oxy2, ost, oer = oracle.orc_klt(a, b, bxy, **mode)
        assert np.array_equal(st, ost) and np.array_equal(nxy, oxy2) and np.array_equal(er, oer)
    # fused chains: trackFrames on the bucketed points, extractStereoFeatures with y_threshold 0.9
    m = int(take(np.int32, 1)[0])
    idx = take(np.int32, m)
    cur = take(np.float32, 2 * m).reshape(m, 2)
    oxy2, ost, _ = oracle.orc_klt(a, b, bxy, win=(21, 21), max_level=3, max_count=50, epsilon=0.001, flags=8)
    keep = np.nonzero(ost)[0]
    assert np.array_equal(idx, keep) and np.array_equal(cur, oxy2[keep])
    m = int(take(np.int32, 1)[0])
    lp = take(np.float32, 2 * m).reshape(m, 2)
    rp = take(np.float32, 2 * m).reshape(m, 2)
    rxy, rst, _ = oracle.orc_klt(a, b, oxy, win=(11, 11), max_level=3, max_count=30, epsilon=0.001, flags=0)
    keep = np.nonzero((rst != 0) & (np.abs(rxy[:, 1] - oxy[:, 1]) < np.float32(0.9)))[0]
    assert m == len(keep) and np.array_equal(lp, oxy[keep]) and np.array_equal(rp, rxy[keep])
    assert int(take(np.int32, 1)[0]) == 0                    # empty inputs -> empty outputs, like OpenCV
    # colour frames: device-side cvtColor(BGR2GRAY) of (a, b, 255 - a), and detect() fed the CV_8UC3 frame directly
    gray = take(np.uint8, w * h).reshape(h, w)
    assert np.array_equal(gray, oracle.orc_bgr2gray(np.stack([a, b, 255 - a], axis=-1)))
    assert int(take(np.int32, 1)[0]) == 1
    assert int(take(np.int32, 1)[0]) == len(oracle.orc_fast(gray, 20, True)[0])
    assert off == len(buf)


@pytest.mark.gpu
def test_frame_cache_uploads_and_builds_each_frame_once(exe, tmp_path):
    """The wrappers' content-addressed frame cache (svo_b200/host/feature.cpp): through the reference-shaped loop
    (LK prev->cur, detect(cur, mask), bucketing) every frame is uploaded once and its pyramid built once, although it is an
    argument of three calls (next of LK, detect, prev of the following LK: src/tracking.cpp:160-165, :82); a repeated LK call
    on the same pair -- also with the stereo window -- launches the LK kernel and nothing else."""
    import json
    w, h, ring, n = 320, 240, 4, 6
    seq = synth.sequence(w, h, 9, ring)
    fin = tmp_path / "frames.bin"
    with open(fin, "wb") as f:
        np.array([w, h, ring, 300, 32, 5, 20], np.int32).tofile(f)
        seq.tofile(f)
    r = subprocess.run([exe, "--bench", str(fin), str(n)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    d = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][0])
    total = n + 10                                         # warm-up frames included in the cache counters
    for mode in ("host_mask", "device_mask"):
        c = d[mode]["cache"]
        assert c["uploads"] == total and c["pyramids"] == total and c["hits"] == 2 * (total - 1), (mode, c)
        assert d[mode]["tracks_per_frame"] > 100 and d[mode]["status_ok_frac"] > 0.9
    assert d["host_mask"]["keypoints_per_frame"] == d["device_mask"]["keypoints_per_frame"]
    rep = d["repeat_call"]
    assert rep == {"launches": 2, "uploads": 0, "pyramids": 0, "hits": 4}, rep


@pytest.mark.gpu
def test_sequence_batch_reference_track_policy(exe, oracle, tmp_path):
    """svo::SequenceBatch with FrontendParams::carryTracks (the reference's loop: survivors carry over, detection on
    key-frames only, src/tracking.cpp:69-70, :168-177, :254) against the CPU pipeline: tracks, key-frame flags and survivor
    indices per sequence and frame, bit-exact."""
    from oracle.cpu_pipeline import CpuFrontend
    from svo_b200.frontend import make_params
    w, h, S, T, max_tracks, ftt = 320, 240, 3, 7, 400, 400
    seqs = np.stack([synth.sequence(w, h, seed, T, times=[v * t for t in range(T)]) for seed, v in ((5, 3.0), (6, 0.1), (7, 1.5))])
    fin, fout = tmp_path / "bin.bin", tmp_path / "bout.bin"
    with open(fin, "wb") as f:
        np.array([w, h, S, T, max_tracks], np.int32).tofile(f)
        np.ascontiguousarray(seqs.transpose(1, 0, 2, 3)).tofile(f)
    r = subprocess.run([exe, "--batch", str(fin), str(fout), str(ftt)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    buf = open(fout, "rb").read()
    off = 0

    def take(dtype, n):
        nonlocal off
        out = np.frombuffer(buf, dtype, n, off)
        off += out.nbytes
        return out

    _, pdict = make_params(w, h, max_tracks=max_tracks, mask_half=10.0, carry_tracks=1, features_to_track=ftt)
    cpus = [CpuFrontend(pdict, impl="orc") for _ in range(S)]
    kfs = []
    for t in range(T):
        for s in range(S):
            ref = cpus[s].step(seqs[s, t])
            nt = int(take(np.int32, 1)[0])
            xy, st, er = take(np.float32, 2 * nt).reshape(nt, 2), take(np.uint8, nt), take(np.float32, nt)
            nb = int(take(np.int32, 1)[0])
            feats = take(np.float32, 2 * nb).reshape(nb, 2)
            nk = int(take(np.int32, 1)[0])
            kf = int(take(np.int32, 1)[0])
            ns = int(take(np.int32, 1)[0])
            surv = take(np.int32, ns)
            assert nt == len(ref["next_xy"]) and np.array_equal(xy, ref["next_xy"]) and np.array_equal(st, ref["status"]), (t, s)
            assert np.array_equal(er, ref["err"]) and kf == ref["keyframe"] and nk == len(ref["kp_xy"]), (t, s)
            assert np.array_equal(feats, ref["bucket_xy"]) and np.array_equal(surv, ref.get("survivor_index", np.zeros(0, np.int32))), (t, s)
            kfs.append(kf)
    assert any(k == 0 for k in kfs[S:]) and any(k == 1 for k in kfs[S:])


@pytest.mark.gpu
def test_frame_cache_notices_foreign_uploads(exe):
    """The wrappers' frame cache tags slots of a context that svo::defaultContext() also hands to the caller.  After direct
    svo_gpu_upload calls into those slots, detect() on a frame the cache "knew" must upload it again (slot generation
    counter, svo_gpu_slot_generation) and return that frame's keypoints, not the foreign image's."""
    r = subprocess.run([exe, "--stale-slot"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "-> ok" in r.stdout
