"""BASELINE.json configs C2-C4 at their FULL sizes (seeds, thresholds, bucket parameters, window and
pyramid depth), through the C ABI.

Bars: FAST and bucketing bit-exact (vs the C oracle, the reference's bucket.cpp and, when importable, OpenCV);
LK bit-exact vs the oracle on a bounded subset of the tracks (the scalar oracle is slow at 31x31 x 6 levels), within
0.01 px / 99.9 % status of cv::calcOpticalFlowPyrLK on ALL tracks, plus the size-independent properties of the domain:
identical pair -> next == prev, translated pair -> translation recovered, pyramid levels idempotent with the oracle.
"""
import numpy as np
import pytest

from svo_b200 import capi, synth

pytestmark = pytest.mark.gpu

LK_TOL_PX = 0.01          # north_star tolerance vs OpenCV
LK_STATUS_AGREE = 0.999

CONFIGS = {
    # name: (w, h, seed, fast thr, bucket size, per bucket, tracks, window, max_level)
    "C2_euroc": (752, 480, 12, 10, 24, 8, 5000, (21, 21), 3),
    "C3_hd": (1920, 1080, 13, 20, 32, 5, 10000, (21, 21), 4),
    "C4_4k": (3840, 2160, 14, 20, 24, 8, 50000, (31, 31), 5),
}


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(device=0, max_w=3840, max_h=2160, max_levels=6, max_points=400000, max_batch=1)
    yield c
    c.close()


@pytest.fixture(scope="module", params=list(CONFIGS))
def cfg(request):
    w, h, seed, thr, bs, k, n, win, ml = CONFIGS[request.param]
    canvas = synth.make_canvas(w, h, seed)
    f0, f1 = synth.render_frame(canvas, w, h, 0), synth.render_frame(canvas, w, h, 1)
    return dict(name=request.param, w=w, h=h, thr=thr, bs=bs, k=k, n=n, win=win, ml=ml, f0=f0, f1=f1)


def test_config_frontend(ctx, oracle, cfg):
    w, h, win, ml = cfg["w"], cfg["h"], cfg["win"], cfg["ml"]
    f0, f1 = cfg["f0"], cfg["f1"]
    ctx.upload(0, f0)
    ctx.upload(1, f1)
    top = ctx.build_pyramid(0, win, ml)      # effective maxLevel, as cv::buildOpticalFlowPyramid returns it
    assert top == ml, (top, ml)
    levels = top + 1
    ctx.build_pyramid(1, win, ml)

    # ---- pyramid: every level bit-exact
    ref = oracle.orc_pyramid(f0, win, ml)
    for l in range(levels):
        assert np.array_equal(ctx.download_level(0, l), ref[l]), (cfg["name"], l)

    # ---- FAST: bit-exact list (positions, order, response)
    xy, resp = ctx.fast(0, cfg["thr"], True)
    oxy, oresp = oracle.orc_fast(f0, cfg["thr"], True)
    assert np.array_equal(xy, oxy) and np.array_equal(resp, oresp), cfg["name"]
    if oracle.have_cv2():
        cxy, cresp = oracle.cv2_fast(f0, cfg["thr"], True)
        assert np.array_equal(xy, cxy) and np.array_equal(resp, cresp), cfg["name"]

    # ---- bucketing: bit-exact (points, ages, order, duplicates) vs the oracle and the reference's own bucket.cpp
    ages = (np.arange(len(xy)) % 7).astype(np.int32)
    bxy, bag = ctx.bucket(xy, ages, w, h, cfg["bs"], cfg["k"])
    oxy2, oag = oracle.orc_bucket(xy, ages, w, h, cfg["bs"], cfg["k"])
    assert np.array_equal(bxy, oxy2) and np.array_equal(bag, oag), cfg["name"]
    if oracle.have_ref_bucket():
        rxy, rag = oracle.ref_bucket(xy, ages, w, h, cfg["bs"], cfg["k"])
        assert np.array_equal(bxy, rxy) and np.array_equal(bag, rag), cfg["name"]
    assert len(bxy) >= cfg["n"], (cfg["name"], len(bxy))
    pts = np.ascontiguousarray(bxy[:cfg["n"]])

    # ---- LK: all tracks vs OpenCV (tolerance), a subset vs the oracle (bit-exact)
    for mode in (dict(win=win, max_level=ml, max_count=50, epsilon=0.001, flags=capi.LK_GET_MIN_EIGENVALS),
                 dict(win=win, max_level=ml, max_count=30, epsilon=0.001, flags=0)):
        nxy, st, er = ctx.klt(0, 1, pts, **mode)
        assert st.mean() > 0.99, (cfg["name"], st.mean())
        sub = np.linspace(0, len(pts) - 1, 1500).astype(np.int64)
        # include the tracks nearest to the image border (they take the zero-padded-derivative path at coarse levels)
        edge = np.argsort(np.minimum.reduce([pts[:, 0], pts[:, 1], w - 1 - pts[:, 0], h - 1 - pts[:, 1]]))[:300]
        sub = np.unique(np.concatenate([sub, edge]))
        oxy3, ost, oer = oracle.orc_klt(f0, f1, np.ascontiguousarray(pts[sub]), **mode)
        assert np.array_equal(st[sub], ost), cfg["name"]
        assert np.array_equal(nxy[sub], oxy3), cfg["name"]
        assert np.array_equal(er[sub], oer), cfg["name"]
        if oracle.have_cv2():
            cxy, cst, cer = oracle.cv2_klt(f0, f1, pts, **mode)
            assert (st == cst).mean() >= LK_STATUS_AGREE, (cfg["name"], (st == cst).mean())
            both = (st == 1) & (cst == 1)
            d = np.abs(nxy[both] - cxy[both]).max()
            assert d <= LK_TOL_PX, (cfg["name"], d)


def test_config_lk_properties(ctx, cfg):
    """Size-independent properties at full size and full track count."""
    w, h, win, ml, n = cfg["w"], cfg["h"], cfg["win"], cfg["ml"], cfg["n"]
    f0 = cfg["f0"]
    mode = dict(win=win, max_level=ml, max_count=50, epsilon=0.001, flags=capi.LK_GET_MIN_EIGENVALS)
    rng = np.random.default_rng(5)
    pts = np.stack([rng.uniform(40, w - 40, n), rng.uniform(40, h - 40, n)], axis=1).astype(np.float32)

    # identical pair: delta = 0 at the first iteration of every level -> next == prev, up to the float32 rounding of
    # (p 2^-L - halfWin) + halfWin at coarse levels (OpenCV's own operation order; exact for most points)
    ctx.upload(0, f0)
    ctx.upload(1, f0)
    ctx.build_pyramid(0, win, ml)
    ctx.build_pyramid(1, win, ml)
    nxy, st, er = ctx.klt(0, 1, pts, **mode)
    assert st.mean() > 0.99
    assert np.abs(nxy[st == 1] - pts[st == 1]).max() < 2e-3
    assert (nxy[st == 1] == pts[st == 1]).mean() > 0.9
    assert (er[st == 1] >= 1e-4).all()

    # integer translation of the frame: recovered to < 0.01 px away from the border
    dx, dy = 3, -2
    g = np.roll(f0, (dy, dx), axis=(0, 1))
    ctx.upload(1, g)
    ctx.build_pyramid(1, win, ml)
    nxy, st, er = ctx.klt(0, 1, pts, **mode)
    ok = st == 1
    assert ok.mean() > 0.99
    d = nxy[ok] - pts[ok] - np.array([dx, dy], np.float32)
    assert np.percentile(np.abs(d), 99.5) < LK_TOL_PX, np.percentile(np.abs(d), 99.5)

    # linearity of the track list: tracking a permutation of the points gives the permuted result (no cross-talk)
    perm = rng.permutation(n)
    nxy2, st2, er2 = ctx.klt(0, 1, np.ascontiguousarray(pts[perm]), **mode)
    assert np.array_equal(nxy2, nxy[perm]) and np.array_equal(st2, st[perm]) and np.array_equal(er2, er[perm])
