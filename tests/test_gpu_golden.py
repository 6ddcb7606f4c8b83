"""GPU parity against the committed golden fixtures (tests/golden/*.npz = outputs of OpenCV 4.13.0 and of the reference's
src/bucket.cpp compiled unmodified; generator tests/golden/make_golden.py).  Nothing here needs cv2 or /root/reference on
the GPU box.  Everything goes through the C ABI (svo_b200.capi -> libsvo_gpu.so)."""
import os

import numpy as np
import pytest

from svo_b200 import capi

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LK_TOL_PX = 0.01          # BASELINE.json north_star
LK_STATUS_AGREE = 0.999


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(device=0, max_w=256, max_h=160, max_levels=5, max_points=40000, max_batch=1)
    yield c
    c.close()


def test_fast_golden(ctx):
    g = np.load(os.path.join(GOLD, "fast.npz"))
    for key in g["cases"]:
        name, t, n = str(key).rsplit("_", 2)
        ctx.upload(0, g["img_" + name])
        xy, resp = ctx.fast(0, int(t[1:]), n == "n1")
        assert np.array_equal(xy, g["xy_" + key].astype(np.float32)), key
        assert np.array_equal(resp, g["resp_" + key].astype(np.float32)), key
    ctx.upload(0, g["img_synth"])
    xy, resp = ctx.fast(0, 20, True, mask=g["mask"])
    assert np.array_equal(xy, g["xy_masked"].astype(np.float32)) and np.array_equal(resp, g["resp_masked"].astype(np.float32))
    xy, resp = ctx.fast_masked_by_points(0, g["mask_pts"], 10.0, 20, True)     # src/tracking.cpp:76-79 on the device
    assert np.array_equal(xy, g["xy_masked"].astype(np.float32)) and np.array_equal(resp, g["resp_masked"].astype(np.float32))


def test_pyramid_golden(ctx):
    g = np.load(os.path.join(GOLD, "pyramid.npz"))
    ctx.upload(1, g["img"])
    assert ctx.build_pyramid(1, (21, 21), 3) == int(g["nlevels"]) - 1
    for l in range(int(g["nlevels"])):
        assert np.array_equal(ctx.download_level(1, l), g["level%d" % l]), l
    ctx.upload(1, g["img2"])
    assert ctx.build_pyramid(1, (21, 21), 4) == int(g["maxlevel2"])            # construction stops early, like OpenCV


@pytest.mark.parametrize("mode", ["temporal", "stereo"])
def test_klt_golden(ctx, mode):
    g = np.load(os.path.join(GOLD, "klt.npz"))
    kw = dict(temporal=dict(win=(21, 21), max_level=3, max_count=50, epsilon=0.001, flags=capi.LK_GET_MIN_EIGENVALS),
              stereo=dict(win=(11, 11), max_level=3, max_count=30, epsilon=0.001, flags=0))[mode]
    ctx.upload(0, g["prev"])
    ctx.upload(1, g["next"])
    ctx.build_pyramid(0, kw["win"], kw["max_level"])
    ctx.build_pyramid(1, kw["win"], kw["max_level"])
    nxy, st, er = ctx.klt(0, 1, g["pts"], **kw)
    cst = g["status_" + mode]
    assert (st == cst).mean() >= LK_STATUS_AGREE
    both = (st == 1) & (cst == 1)
    assert both.sum() > 100
    assert np.abs(nxy[both] - g["next_" + mode][both]).max() <= LK_TOL_PX
    assert np.allclose(er[both], g["err_" + mode][both], rtol=5e-3, atol=1e-5)


def test_bucket_golden(ctx):
    g = np.load(os.path.join(GOLD, "bucket.npz"))
    for name in g["cases"]:
        w, h, bs, k = (int(v) for v in g["cfg_" + name])
        gx, ga = ctx.bucket(g["in_xy_" + name], g["in_age_" + name], w, h, bs, k)
        assert np.array_equal(gx, g["out_xy_" + name]) and np.array_equal(ga, g["out_age_" + name]), name
