"""Seeded random sweeps of the C ABI against the CPU oracle: ragged image sizes (down to 1x1), every alignment of the row
stride, random thresholds / masks / windows / criteria, empty and oversize point lists, points outside the image.  Same
bars as tests/test_gpu_parity.py: FAST, mask, pyramid and bucketing bit-exact; LK bit-exact against the oracle's
exact-integer accumulation (which is within 1e-3 px of OpenCV, tests/test_oracle_pin.py)."""
import numpy as np
import pytest

from svo_b200 import capi, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(device=0, max_w=400, max_h=300, max_levels=5, max_points=120000, max_batch=1)
    yield c
    c.close()


def _texture(rng, w, h):
    """A frame with corners at a sane density: blurred noise, noise, blocks or a flat image, picked at random."""
    kind = rng.integers(0, 4)
    if kind == 0 and w >= 8 and h >= 8:
        c = synth.make_canvas(w, h, int(rng.integers(1, 1 << 30)))
        return synth.render_frame(c, w, h, float(rng.uniform(0, 3)))
    if kind == 1:
        return synth.block_image(w, h, int(rng.integers(1, 1 << 30)), block=int(rng.integers(2, 9)))
    if kind == 2:
        return np.full((h, w), int(rng.integers(0, 256)), np.uint8)
    return rng.integers(0, 256, size=(h, w), dtype=np.uint8)


def test_fuzz_fast_and_mask(ctx, oracle):
    rng = np.random.default_rng(2024)
    for case in range(60):
        w, h = int(rng.integers(1, 300)), int(rng.integers(1, 200))
        img = _texture(rng, w, h)
        if rng.integers(0, 2):                                    # a cv::Mat ROI: rows strided, any alignment
            pad = int(rng.integers(1, 9))
            big = np.zeros((h, w + pad + 3), np.uint8)
            off = int(rng.integers(0, 4))
            big[:, off:off + w] = img
            img = big[:, off:off + w]
        thr = int(rng.choice([0, 1, 5, 20, 60, 200, 255, 300, -3]))
        nms = bool(rng.integers(0, 2))
        mask = None
        if rng.integers(0, 2):
            mask = (rng.random((h, w)) < rng.uniform(0, 1)).astype(np.uint8) * int(rng.integers(1, 256))
        ctx.upload(0, img)
        gx, gr = ctx.fast(0, thr, nms, mask=mask)
        ox, orr = oracle.orc_fast(np.ascontiguousarray(img), thr, nms, mask=mask)
        assert np.array_equal(gx, ox) and np.array_equal(gr, orr), (case, w, h, thr, nms)
        n = int(rng.integers(0, 40))
        pts = rng.uniform([-20, -20], [w + 20, h + 20], (n, 2)).astype(np.float32)
        pts[: n // 2] = np.round(pts[: n // 2] * 2) / 2           # .5 coordinates: round-half-even of the box corners
        half = float(rng.choice([10.0, 0.5, 3.0, 25.5]))
        gx, gr = ctx.fast_masked_by_points(0, pts, half, thr, nms)
        ox, orr = oracle.orc_fast(np.ascontiguousarray(img), thr, nms, mask=oracle.orc_build_mask(w, h, pts, half))
        assert np.array_equal(gx, ox) and np.array_equal(gr, orr), (case, "points mask", w, h, half)


def test_fuzz_pyramid(ctx, oracle):
    rng = np.random.default_rng(7)
    for case in range(40):
        w, h = int(rng.integers(1, 400)), int(rng.integers(1, 300))
        win = (int(rng.integers(3, 34)), int(rng.integers(3, 34)))
        ml = int(rng.integers(0, 5))
        img = rng.integers(0, 256, size=(h, w), dtype=np.uint8)
        ctx.upload(0, img)
        lv = ctx.build_pyramid(0, win, ml)
        want = oracle.orc_pyramid(img, win, ml)
        assert lv == len(want) - 1, (case, w, h, win, ml)
        for l, p in enumerate(want):
            assert np.array_equal(ctx.download_level(0, l), p), (case, w, h, win, ml, l)


@pytest.mark.parametrize("mode", ["1", "2"])
def test_fuzz_bucket(oracle, monkeypatch, mode):
    monkeypatch.setenv("SVO_GPU_BUCKET_MODE", mode)
    rng = np.random.default_rng(99)
    with capi.Context(device=0, max_w=2000, max_h=1200, max_levels=2, max_points=60000, max_batch=1) as c:
        for case in range(40):
            w, h = int(rng.integers(20, 2000)), int(rng.integers(20, 1200))
            bs = int(rng.integers(3, 120))
            k = int(rng.integers(1, 9))
            n = int(rng.choice([0, 1, 2, 17, 300, 5000, 30000]))
            xy = rng.uniform([0, 0], [w, h], (n, 2)).astype(np.float32)
            if rng.integers(0, 2):
                xy = np.floor(xy)                                  # FAST-like integer coordinates, cell borders included
            if n and rng.integers(0, 2):
                xy = xy[np.lexsort((xy[:, 0], xy[:, 1]))]          # raster order, as FAST emits
            ages = rng.integers(0, 100, n).astype(np.int32)
            gx, ga = c.bucket(xy, ages, w, h, bs, k)
            ox, oa = oracle.orc_bucket(xy, ages, w, h, bs, k)
            assert np.array_equal(gx, ox) and np.array_equal(ga, oa), (case, w, h, bs, k, n)


def test_fuzz_klt(ctx, oracle):
    rng = np.random.default_rng(31)
    for case in range(36):
        w, h = int(rng.integers(24, 330)), int(rng.integers(24, 250))
        canvas = synth.make_canvas(w, h, int(rng.integers(1, 1 << 30)))
        a = synth.render_frame(canvas, w, h, 0.0)
        b = synth.render_frame(canvas, w, h, float(rng.uniform(0.2, 2.5)))
        if case % 6 == 5:
            b = rng.integers(0, 256, size=(h, w), dtype=np.uint8)      # nothing to track: early exits, status 0
        win = [(21, 21), (11, 11), (31, 31), (int(rng.integers(3, 33)), int(rng.integers(3, 33)))][case % 4]
        mode = dict(win=win, max_level=int(rng.integers(0, 5)), max_count=int(rng.choice([1, 5, 30, 50, 150])),
                    epsilon=float(rng.choice([0.0, 0.001, 0.01, 0.3, 20.0])),
                    criteria_type=int(rng.choice([1, 2, 3])), flags=int(rng.choice([0, capi.LK_GET_MIN_EIGENVALS])),
                    min_eig_threshold=float(rng.choice([1e-4, 1e-2, 0.0])))
        n = int(rng.choice([0, 1, 33, 200]))
        pts = rng.uniform([-40, -40], [w + 40, h + 40], (n, 2)).astype(np.float32)
        pts[: n // 3] = np.floor(pts[: n // 3])                         # zero fractional parts: weights (16384, 0, 0, 0)
        ctx.upload(0, a)
        ctx.upload(1, b)
        ctx.build_pyramid(0, win, mode["max_level"])
        ctx.build_pyramid(1, win, mode["max_level"])
        nxy, st, er = ctx.klt(0, 1, pts, **mode)
        oxy, ost, oer, iters = oracle.orc_klt(a, b, pts, return_iters=True, **mode)
        tag = (case, w, h, mode, n)
        assert np.array_equal(st, ost), tag
        assert np.array_equal(nxy, oxy), tag
        assert np.array_equal(er, oer), tag
        assert ctx.last_klt_iterations == iters, tag
