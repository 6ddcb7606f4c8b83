"""The filters that follow each LK call in the reference and the two fused call chains, through
the C ABI, against the CPU oracle run stage by stage exactly as src/tracking.cpp:74-122 and :154-179 do."""
import numpy as np
import pytest

from svo_b200 import capi

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(device=0, max_w=1280, max_h=512, max_levels=4, max_points=60000, max_batch=1)
    yield c
    c.close()


def _ref_filter(a, b, st, y_thr):
    keep = st != 0
    if y_thr >= 0:
        keep &= np.abs(b[:, 1] - a[:, 1]) < np.float32(y_thr)      # float32 arithmetic, src/tracking.cpp:110
    idx = np.nonzero(keep)[0].astype(np.int32)
    return idx, a[idx], b[idx]


@pytest.mark.parametrize("n", [0, 1, 31, 1024, 1025, 2000, 50000])
@pytest.mark.parametrize("y_thr", [-1.0, 0.0, 0.75, 40.0])
def test_filter_tracks(ctx, n, y_thr):
    rng = np.random.default_rng(n + 7)
    a = rng.uniform(0, 1000, (n, 2)).astype(np.float32)
    b = (a + rng.normal(0, 1.0, (n, 2))).astype(np.float32)
    if n > 3:
        b[3, 1] = a[3, 1] + np.float32(0.75)       # exactly on the threshold: strict `<` rejects
        b[2, 1] = np.nan                           # NaN never passes `<`
    st = (rng.uniform(size=n) < 0.8).astype(np.uint8)
    idx, ao, bo = ctx.filter_tracks(a, b, st, y_thr)
    ridx, ra, rb = _ref_filter(a, b, st, y_thr)
    assert np.array_equal(idx, ridx)
    assert np.array_equal(ao, ra) and np.array_equal(bo, rb, equal_nan=True)


def test_filter_all_and_none(ctx):
    a = np.zeros((3000, 2), np.float32)
    idx, _, _ = ctx.filter_tracks(a, a, np.ones(3000, np.uint8))
    assert np.array_equal(idx, np.arange(3000))
    idx, _, _ = ctx.filter_tracks(a, a, np.zeros(3000, np.uint8))
    assert len(idx) == 0


def test_track_frames_matches_reference_loop(ctx, oracle, c1_frames):
    """Tracking::trackFrames: LK 21x21 then `if (status[i])` compaction (src/tracking.cpp:154-179)."""
    a, b = c1_frames
    kxy, _ = oracle.orc_fast(a, 20, True)
    pts, _ = oracle.orc_bucket(kxy, np.zeros(len(kxy), np.int32), a.shape[1], a.shape[0], 32, 5)
    pts = np.concatenate([pts[:2000], np.array([[0.0, 0.0], [-50.0, 10.0], [1240.0, 375.0], [3000.0, 3000.0]], np.float32)])
    ctx.upload(0, a)
    ctx.upload(1, b)
    ctx.build_pyramid(0, (21, 21), 3)
    ctx.build_pyramid(1, (21, 21), 3)
    mode = dict(win=(21, 21), max_level=3, max_count=50, epsilon=0.001, flags=capi.LK_GET_MIN_EIGENVALS)
    idx, cur, er = ctx.track_frames(0, 1, pts, **mode)
    nxy, st, oer = ctx.klt(0, 1, pts, **mode)
    keep = np.nonzero(st)[0]
    assert 0 < len(keep) < len(pts)                # the far-outside points are lost
    assert np.array_equal(idx, keep) and np.array_equal(cur, nxy[keep]) and np.array_equal(er, oer[keep])
    sub = np.concatenate([np.arange(0, 2000, 7), np.arange(2000, 2004)])
    oxy, ost, _ = oracle.orc_klt(a, b, np.ascontiguousarray(pts[sub]), **mode)
    sel = np.isin(sub, idx)
    assert np.array_equal(ost.astype(bool), sel)
    assert np.array_equal(oxy[sel], cur[np.searchsorted(idx, sub[sel])])


@pytest.mark.parametrize("bucket", [(0, 0), (32, 5)])
@pytest.mark.parametrize("y_thr", [40.0, 0.9, -1.0])
def test_keyframe_matches_reference_chain(ctx, oracle, small_frames, bucket, y_thr):
    """extractFeatures + findLeftFeaturesInRight (src/tracking.cpp:74-122), every stage bit-exact vs the oracle."""
    left, right = small_frames
    h, w = left.shape
    rng = np.random.default_rng(3)
    prev = np.stack([rng.uniform(-5, w + 5, 150), rng.uniform(-5, h + 5, 150)], axis=1).astype(np.float32)
    ctx.upload(0, left)
    ctx.upload(1, right)
    ctx.build_pyramid(0, (21, 21), 3)        # one pyramid serves the 21x21 temporal and the 11x11 stereo call
    ctx.build_pyramid(1, (21, 21), 3)
    lo, ro, nd = ctx.keyframe(0, 1, prev, fast_threshold=20, fast_nms=True, mask_half=10.0, bucket_size=bucket[0],
                              per_bucket=bucket[1], y_threshold=y_thr)
    mask = oracle.orc_build_mask(w, h, prev, 10.0)
    kxy, _ = oracle.orc_fast(left, 20, True, mask)
    if bucket[0] > 0:
        kxy, _ = oracle.orc_bucket(kxy, np.zeros(len(kxy), np.int32), w, h, bucket[0], bucket[1])
    assert nd == len(kxy) and nd > 50
    rxy, st, _ = oracle.orc_klt(left, right, kxy, win=(11, 11), max_level=3, max_count=30, epsilon=0.001, flags=0)
    _, ra, rb = _ref_filter(kxy, rxy, st, y_thr)
    assert np.array_equal(lo, ra) and np.array_equal(ro, rb)
    if y_thr == 0.9:
        assert 0 < len(lo) < int(st.sum())     # the epipolar test removed some tracked points


def test_keyframe_without_previous_features(ctx, oracle, small_frames):
    left, right = small_frames
    ctx.upload(0, left)
    ctx.upload(1, right)
    ctx.build_pyramid(0, (11, 11), 3)
    ctx.build_pyramid(1, (11, 11), 3)
    lo, ro, nd = ctx.keyframe(0, 1, np.zeros((0, 2), np.float32))
    kxy, _ = oracle.orc_fast(left, 20, True)
    assert nd == len(kxy)
    rxy, st, _ = oracle.orc_klt(left, right, kxy, win=(11, 11), max_level=3, max_count=30, epsilon=0.001, flags=0)
    _, ra, rb = _ref_filter(kxy, rxy, st, 40.0)
    assert np.array_equal(lo, ra) and np.array_equal(ro, rb)
