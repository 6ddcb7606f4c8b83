"""GPU parity tests: the sm_100a kernels, called through the C ABI (svo_b200.capi -> libsvo_gpu.so), against the CPU
oracle (oracle/svo_oracle.c), the reference's own bucket.cpp (oracle/_ref) and, when importable, OpenCV itself.

Bars (BASELINE.json north_star): FAST detections and bucket selections bit-exact (positions, order, responses, ages);
LK within 0.01 px of cv::calcOpticalFlowPyrLK with >= 99.9 % status agreement -- and bit-exact against the oracle's
exact-integer accumulation mode, which is what the kernel implements.
"""
import numpy as np
import pytest

from svo_b200 import capi, synth

pytestmark = pytest.mark.gpu

LK_TOL_PX = 0.01          # north_star tolerance vs OpenCV
LK_STATUS_AGREE = 0.999


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(device=0, max_w=1280, max_h=512, max_levels=5, max_points=480000, max_batch=2)
    yield c
    c.close()


def _reflect_pad(a, m):
    """BORDER_REFLECT_101 by m pixels, also when m exceeds the image size (index map applied repeatedly, as OpenCV does)."""
    def idx(n):
        i = np.arange(-m, n + m)
        if n == 1:
            return np.zeros_like(i)
        period = 2 * n - 2
        i = np.mod(i, period)
        return np.where(i >= n, period - i, i)
    return a[np.ix_(idx(a.shape[0]), idx(a.shape[1]))]


# ------------------------------------------------------------------------------------------------ FAST
def _images():
    canvas = synth.make_canvas(1241, 376, 11)
    yield "c1", synth.render_frame(canvas, 1241, 376, 0)
    yield "noise", synth.noise_image(1241, 376, 1)
    yield "blocks", synth.block_image(1241, 376, 2)
    yield "odd", synth.noise_image(333, 77, 3)
    yield "tiny7", synth.noise_image(7, 7, 4)
    yield "tiny8", synth.noise_image(8, 8, 5)
    yield "thin", synth.noise_image(40, 6, 6)
    yield "w65", synth.noise_image(65, 17, 7)


@pytest.mark.parametrize("thr", [20, 7, 0, 100])
@pytest.mark.parametrize("nms", [True, False])
def test_fast_bit_exact_vs_oracle(ctx, oracle, thr, nms):
    for name, img in _images():
        ctx.upload(0, img)
        xy, resp = ctx.fast(0, thr, nms)
        oxy, oresp = oracle.orc_fast(img, thr, nms)
        assert len(xy) == len(oxy), (name, thr, nms, len(xy), len(oxy))
        assert np.array_equal(xy, oxy), (name, thr, nms)
        assert np.array_equal(resp, oresp), (name, thr, nms)


def test_fast_vs_cv2(ctx, oracle):
    if not oracle.have_cv2():
        pytest.skip("cv2 not importable on this box")
    for name, img in _images():
        for thr in (20, 10):
            ctx.upload(0, img)
            xy, resp = ctx.fast(0, thr, True)
            cxy, cresp = oracle.cv2_fast(img, thr, True)
            assert np.array_equal(xy, cxy) and np.array_equal(resp, cresp), (name, thr)


def test_fast_mask(ctx, oracle, c1_frames):
    img = c1_frames[0]
    h, w = img.shape
    pts = np.random.default_rng(3).uniform([-5, -5], [w + 5, h + 5], size=(500, 2)).astype(np.float32)
    pts[:10] = np.round(pts[:10]) + 0.5          # exercise round-half-even
    mask = oracle.orc_build_mask(w, h, pts, 10.0)
    ctx.upload(0, img)
    xy, resp = ctx.fast(0, 20, True, mask=mask)
    oxy, oresp = oracle.orc_fast(img, 20, True, mask=mask)
    assert len(oxy) > 100
    assert np.array_equal(xy, oxy) and np.array_equal(resp, oresp)
    # the device-built mask (src/tracking.cpp:76-79) gives the same list
    xy2, resp2 = ctx.fast_masked_by_points(0, pts, 10.0, 20, True)
    assert np.array_equal(xy2, oxy) and np.array_equal(resp2, oresp)
    # strided host mask (a cv::Mat ROI)
    big = np.zeros((h, w + 13), np.uint8)
    big[:, :w] = mask
    xy3, _ = ctx.fast(0, 20, True, mask=big[:, :w])
    assert np.array_equal(xy3, oxy)


@pytest.mark.parametrize("cta", ["1", "0", "2"])
def test_fast_cta_sizes(oracle, c1_frames, monkeypatch, cta):
    """SVO_GPU_FAST_SMALL forces the FAST CTA size (1 = 128 threads, 0 = 256, 2 = 512; unset = by tile count): same list."""
    img = c1_frames[0]
    h, w = img.shape
    pts = np.random.default_rng(4).uniform([0, 0], [w, h], size=(800, 2)).astype(np.float32)
    monkeypatch.setenv("SVO_GPU_FAST_SMALL", cta)
    with capi.Context(max_w=w, max_h=h, max_levels=4, max_points=65536, max_batch=1) as c:
        c.upload(0, img)
        xy, resp = c.fast_masked_by_points(0, pts, 10.0, 20, True)
    oxy, oresp = oracle.orc_fast(img, 20, True, mask=oracle.orc_build_mask(w, h, pts, 10.0))
    assert len(oxy) > 1000 and np.array_equal(xy, oxy) and np.array_equal(resp, oresp)


def test_fast_capacity_error(ctx, c1_frames):
    ctx.upload(0, c1_frames[0])
    with pytest.raises(capi.SvoGpuError) as e:
        ctx.fast(0, 20, True, capacity=10)
    assert e.value.code == capi.E_CAPACITY


# ------------------------------------------------------------------------------------------------ pyramid
@pytest.mark.parametrize("shape", [(1241, 376), (333, 77), (64, 48), (101, 203)])
def test_pyramid_bit_exact(ctx, oracle, shape):
    w, h = shape
    img = synth.noise_image(w, h, 9)
    ctx.upload(1, img)
    for win, ml in (((21, 21), 3), ((11, 11), 4), ((31, 31), 2)):
        levels = ctx.build_pyramid(1, win, ml)
        ref = oracle.orc_pyramid(img, win, ml)
        assert levels == len(ref) - 1
        for l in range(levels + 1):
            assert np.array_equal(ctx.download_level(1, l), ref[l]), (shape, win, l)
            padded, m = ctx.download_level_padded(1, l)      # the BORDER_REFLECT_101 margin LK reads
            assert m >= max(win)
            assert np.array_equal(padded, _reflect_pad(ref[l], m)), (shape, win, l)


@pytest.mark.parametrize("shape", [(1241, 376), (130, 65), (64, 48), (1920, 1080)])
def test_pyramid_per_level_launches(oracle, shape, monkeypatch):
    """SVO_GPU_PYR_FUSED=0: ingest + one pyrDown launch per level (the path tiny levels and levels >= 4 always take) gives the
    same padded levels as the fused kernel every other test runs, and as the oracle."""
    w, h = shape
    img = synth.noise_image(w, h, 10)
    got = {}
    for fused in ("1", "0"):
        monkeypatch.setenv("SVO_GPU_PYR_FUSED", fused)
        with capi.Context(max_w=w, max_h=h, max_levels=6, max_points=1024, max_batch=1) as c:
            c.upload(0, img)
            n0 = c.launch_count
            levels = c.build_pyramid(0, (21, 21), 5)
            got[fused] = (levels, c.launch_count - n0, [c.download_level_padded(0, l)[0] for l in range(levels + 1)])
    ref = oracle.orc_pyramid(img, (21, 21), 5)
    assert got["1"][0] == got["0"][0] == len(ref) - 1
    assert got["0"][1] == 2 + got["0"][0]                       # ingest, ingest edge, one launch per further level
    if min(w, h) > 2 * 33:                                       # margin 32 narrower than levels 0 and 1: the fused kernel ran
        assert got["1"][1] < got["0"][1]
    for l, (a, b) in enumerate(zip(got["1"][2], got["0"][2])):
        assert np.array_equal(a, b), (shape, l)
        assert np.array_equal(a, _reflect_pad(ref[l], (a.shape[0] - ref[l].shape[0]) // 2)), (shape, l)


# ------------------------------------------------------------------------------------------------ bucketing
def _bucket_cases():
    rng = np.random.default_rng(0)
    yield "toy", np.array([[1, 1], [2, 2], [3, 3], [4, 4], [5, 5]], np.float32), np.array([5, 1, 9, 0, 7], np.int32), 100, 100, 50, 2
    pts = rng.uniform([0, 0], [1241, 376], size=(5000, 2)).astype(np.float32)
    yield "uniform5000", pts, rng.integers(0, 50, 5000).astype(np.int32), 1241, 376, 50, 4
    yield "k1", pts, np.zeros(5000, np.int32), 1241, 376, 32, 1
    yield "k8", pts[:3000], rng.integers(0, 9, 3000).astype(np.int32), 752, 480, 24, 8
    yield "bigcell", pts[:100], np.arange(100, dtype=np.int32), 1241, 376, 2000, 3
    yield "empty", np.zeros((0, 2), np.float32), np.zeros(0, np.int32), 640, 480, 50, 4
    ipts = np.floor(pts)
    yield "integer", ipts, rng.integers(0, 5, 5000).astype(np.int32), 1241, 376, 32, 5
    dense = rng.uniform([0, 0], [200, 100], size=(20000, 2)).astype(np.float32)
    yield "dense", dense, rng.integers(0, 100, 20000).astype(np.int32), 200, 100, 10, 3


def test_bucket_bit_exact(ctx, oracle):
    for name, xy, ages, w, h, bs, k in _bucket_cases():
        gx, ga = ctx.bucket(xy, ages, w, h, bs, k)
        ox, oa = oracle.orc_bucket(xy, ages, w, h, bs, k)
        assert np.array_equal(gx, ox) and np.array_equal(ga, oa), name
        if oracle.have_ref_bucket():
            rx, ra = oracle.ref_bucket(xy, ages, w, h, bs, k)
            assert np.array_equal(gx, rx) and np.array_equal(ga, ra), name


def test_bucket_on_fast_output(ctx, oracle, c1_frames):
    ctx.upload(0, c1_frames[0])
    xy, _ = ctx.fast(0, 20, True)
    gx, ga = ctx.bucket(xy, np.zeros(len(xy), np.int32), 1241, 376, 32, 5)
    ox, oa = oracle.orc_bucket(xy, np.zeros(len(xy), np.int32), 1241, 376, 32, 5)
    assert len(gx) > 2000 and np.array_equal(gx, ox) and np.array_equal(ga, oa)


def test_bucket_rejects_out_of_table(ctx):
    with pytest.raises(capi.SvoGpuError) as e:
        ctx.bucket(np.array([[-500.0, 10.0]], np.float32), np.zeros(1, np.int32), 640, 480, 50, 4)
    assert e.value.code == capi.E_ARG


# ------------------------------------------------------------------------------------------------ LK
def _lk_points(oracle, img, n, seed=0):
    xy, _ = oracle.orc_fast(img, 20, True)
    rng = np.random.default_rng(seed)
    sel = rng.permutation(len(xy))[:n]
    pts = xy[np.sort(sel)].copy()
    pts += rng.uniform(-0.5, 0.5, pts.shape).astype(np.float32)   # sub-pixel starts
    h, w = img.shape
    adversarial = np.array([[0, 0], [w - 1, h - 1], [w - 0.5, 10], [5, h - 2], [-30, -30], [w + 25, h + 25],
                            [-12.3, 50.2], [w / 2, -9.5], [3.25, 3.75], [w - 3.5, h - 3.5]], np.float32)
    return np.concatenate([pts, adversarial]).astype(np.float32)


MODES = [
    dict(win=(21, 21), max_level=3, max_count=50, epsilon=0.001, flags=capi.LK_GET_MIN_EIGENVALS),   # tracking.cpp:160-165
    dict(win=(11, 11), max_level=3, max_count=30, epsilon=0.001, flags=0),                           # tracking.cpp:101-105
    dict(win=(31, 31), max_level=2, max_count=20, epsilon=0.01, flags=0),
    dict(win=(15, 9), max_level=2, max_count=30, epsilon=0.01, flags=capi.LK_GET_MIN_EIGENVALS),     # generic kernel
    dict(win=(21, 21), max_level=0, max_count=10, epsilon=0.03, flags=0),
]


@pytest.mark.parametrize("mode", MODES)
def test_klt_bit_exact_vs_oracle(ctx, oracle, small_frames, mode):
    a, b = small_frames
    pts = _lk_points(oracle, a, 300)
    ctx.upload(0, a)
    ctx.upload(1, b)
    ctx.build_pyramid(0, mode["win"], mode["max_level"])
    ctx.build_pyramid(1, mode["win"], mode["max_level"])
    nxy, st, er = ctx.klt(0, 1, pts, **mode)
    oxy, ost, oer, iters = oracle.orc_klt(a, b, pts, return_iters=True, **mode)
    assert np.array_equal(st, ost)
    assert np.array_equal(nxy, oxy)
    assert np.array_equal(er, oer)
    assert ctx.last_klt_iterations == iters


@pytest.mark.parametrize("mode", MODES[:3])
def test_klt_vs_cv2(ctx, oracle, c1_frames, mode):
    if not oracle.have_cv2():
        pytest.skip("cv2 not importable on this box")
    a, b = c1_frames
    pts = _lk_points(oracle, a, 2000)
    ctx.upload(0, a)
    ctx.upload(1, b)
    ctx.build_pyramid(0, mode["win"], mode["max_level"])
    ctx.build_pyramid(1, mode["win"], mode["max_level"])
    nxy, st, er = ctx.klt(0, 1, pts, **mode)
    cxy, cst, cer = oracle.cv2_klt(a, b, pts, **mode)
    agree = (st == cst)
    assert agree.mean() >= LK_STATUS_AGREE, agree.mean()
    both = (st == 1) & (cst == 1)
    assert both.sum() > 1500
    d = np.abs(nxy[both] - cxy[both]).max()
    assert d <= LK_TOL_PX, d
    # err is not read by the reference (src/tracking.cpp:159-165); OpenCV accumulates it in float32 SIMD lanes, the kernel
    # exactly in integers, hence a looser bound than for the positions
    assert np.allclose(er[both], cer[both], rtol=5e-3, atol=1e-5)


def test_klt_negative_corner_weight(ctx, oracle, small_frames):
    """Fractions of ~0.6 / 2^14 round iw00, iw01, iw10 all upwards, so iw11 = 2^14 - (their sum) = -1: the bilinear weights
    must be treated as signed (a DP2A with unsigned 16-bit weights would read 65535)."""
    a, b = small_frames
    base = np.floor(_lk_points(oracle, a, 200)[:200])
    pts = (base + np.float32(3.8e-5)).astype(np.float32)
    assert np.all(pts != base)
    for mode in MODES[:2]:
        ctx.upload(0, a)
        ctx.upload(1, b)
        ctx.build_pyramid(0, mode["win"], mode["max_level"])
        ctx.build_pyramid(1, mode["win"], mode["max_level"])
        nxy, st, er = ctx.klt(0, 1, pts, **dict(mode, max_level=0))
        oxy, ost, oer = oracle.orc_klt(a, b, pts, **dict(mode, max_level=0))
        assert np.array_equal(st, ost) and np.array_equal(nxy, oxy) and np.array_equal(er, oer)


def test_klt_known_answers(ctx, oracle, c1_frames):
    a, _ = c1_frames
    pts = _lk_points(oracle, a, 500)[:500]
    mode = MODES[0]
    ctx.upload(0, a)
    ctx.upload(1, a)
    ctx.build_pyramid(0, mode["win"], mode["max_level"])
    ctx.build_pyramid(1, mode["win"], mode["max_level"])
    nxy, st, er = ctx.klt(0, 1, pts, **mode)          # identical pair: delta = 0 at the first iteration
    assert st.all() and np.abs(nxy - pts).max() < 2e-4   # (p - halfWin) + halfWin in float32, as in OpenCV
    ipts = np.round(pts)
    nxy, st, er = ctx.klt(0, 1, ipts, **mode)
    assert st.all() and np.array_equal(nxy, ipts)     # integer starts come back exactly
    const = np.full_like(a, 77)
    ctx.upload(0, const)
    ctx.upload(1, const)
    ctx.build_pyramid(0, mode["win"], mode["max_level"])
    ctx.build_pyramid(1, mode["win"], mode["max_level"])
    nxy, st, er = ctx.klt(0, 1, pts, **mode)          # constant image: minEig = 0 -> all lost
    assert not st.any() and np.all(er == 0) and np.abs(nxy - pts).max() < 2e-4
    nxy, st, er = ctx.klt(0, 1, np.zeros((0, 2), np.float32), **mode)
    assert len(nxy) == 0 and len(st) == 0 and len(er) == 0


def test_klt_unsupported_and_state(ctx, c1_frames):
    ctx.upload(0, c1_frames[0])
    ctx.upload(1, c1_frames[1])
    ctx.build_pyramid(0, (11, 11), 3)
    ctx.build_pyramid(1, (11, 11), 3)
    pts = np.array([[100, 100]], np.float32)
    with pytest.raises(capi.SvoGpuError) as e:
        ctx.klt(0, 1, pts, flags=capi.LK_USE_INITIAL_FLOW)
    assert e.value.code == capi.E_UNSUPPORTED
    with pytest.raises(capi.SvoGpuError) as e:
        ctx.klt(0, 1, pts, win=(41, 41))              # pyramids were built for a smaller window
    assert e.value.code == capi.E_STATE


def test_klt_refuses_pyramids_shallower_than_max_level(ctx, small_frames):
    """cv::calcOpticalFlowPyrLK(maxLevel = 3) on a 320x240 frame uses levels 0..3.  Pyramids built with maxLevel 1 hold two:
    running LK on them would return different tracks than OpenCV with SVO_GPU_OK, so the direct C-ABI calls refuse
    (the C++ wrappers rebuild the pyramid instead, svo_b200/host/feature.cpp:need_pyramid)."""
    a, b = small_frames
    ctx.upload(0, a)
    ctx.upload(1, b)
    ctx.build_pyramid(0, (21, 21), 1)
    ctx.build_pyramid(1, (21, 21), 1)
    pts = np.array([[100.0, 100.0], [50.5, 60.25]], np.float32)
    ctx.klt(0, 1, pts, win=(21, 21), max_level=1, max_count=30, epsilon=0.01)          # what the slots hold: fine
    for call in (lambda: ctx.klt(0, 1, pts, win=(21, 21), max_level=3, max_count=30, epsilon=0.01),
                 lambda: ctx.track_frames(0, 1, pts, win=(21, 21), max_level=3)):
        with pytest.raises(capi.SvoGpuError) as e:
            call()
        assert e.value.code == capi.E_STATE and "levels" in str(e.value)
    ctx.build_pyramid(0, (21, 21), 3)
    ctx.build_pyramid(1, (21, 21), 3)
    ctx.klt(0, 1, pts, win=(21, 21), max_level=3, max_count=30, epsilon=0.01)
    # asking for more levels than the image supports is not an error: OpenCV stops at the same level
    ctx.klt(0, 1, pts, win=(21, 21), max_level=4, max_count=30, epsilon=0.01)
