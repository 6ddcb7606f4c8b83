"""Colour-frame ingest on the device: svo_gpu_upload_bgr* must give cv::cvtColor(img, COLOR_BGR2GRAY)
bit for bit -- the conversion the reference's loader runs on the host (include/async_image_loader.h:66-69) -- against the
committed cv2 fixture (tests/golden/bgr.npz) and against the oracle on ragged widths, strided rows, any source alignment,
batches and caller-owned device memory; and the rest of the path must see exactly that gray image."""
import os

import numpy as np
import pytest

from svo_b200 import capi

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(device=0, max_w=1241, max_h=376, max_levels=4, max_points=40000, max_batch=4)
    yield c
    c.close()


def test_bgr_golden(ctx):
    g = np.load(os.path.join(GOLD, "bgr.npz"))
    for name in ("rnd", "extreme"):
        img = g[name]
        ctx.upload_bgr(0, img)
        assert np.array_equal(ctx.download_image(0, img.shape[1], img.shape[0]), g[name + "_gray"]), name


@pytest.mark.parametrize("w,h", [(1, 1), (3, 2), (4, 5), (5, 3), (127, 9), (128, 4), (129, 7), (1241, 376)])
def test_bgr_vs_oracle_ragged(ctx, oracle, w, h):
    rng = np.random.default_rng(w * 1000 + h)
    img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
    ctx.upload_bgr(0, img)
    assert np.array_equal(ctx.download_image(0, w, h), oracle.orc_bgr2gray(img))


def test_bgr_strided_and_misaligned(ctx, oracle):
    """cv::Mat ROI-style inputs: row stride > 3 w and a first byte at every alignment 0..3."""
    rng = np.random.default_rng(9)
    big = rng.integers(0, 256, size=(40, 3 * 70 + 8), dtype=np.uint8)
    for off in range(4):
        w, h = 61, 37
        view = np.lib.stride_tricks.as_strided(big.reshape(-1)[off:], shape=(h, w, 3), strides=(big.strides[0], 3, 1))
        ctx.upload_bgr(0, view)
        assert np.array_equal(ctx.download_image(0, w, h), oracle.orc_bgr2gray(np.ascontiguousarray(view))), off


def test_bgr_batch_host_and_device(ctx, oracle):
    import torch
    rng = np.random.default_rng(4)
    imgs = rng.integers(0, 256, size=(3, 50, 77, 3), dtype=np.uint8)
    want = [oracle.orc_bgr2gray(im) for im in imgs]
    ctx.upload_bgr_batch(1, imgs)
    for b in range(3):
        assert np.array_equal(ctx.download_image(1, 77, 50, lane=b), want[b]), b
    dev = torch.from_numpy(imgs).cuda()
    ctx.upload_bgr_batch_device(2, dev.data_ptr(), 77, 50, 3 * 77, 3 * 77 * 50, 3)
    for b in range(3):
        assert np.array_equal(ctx.download_image(2, 77, 50, lane=b), want[b]), b


def test_bgr_feeds_the_front_end(ctx, oracle, small_frames):
    """A gray frame replicated into B = G = R converts back to itself (3735 + 19235 + 9798 = 2^15), so FAST and the
    pyramid on a BGR upload must equal those on the gray upload."""
    a, _ = small_frames
    bgr = np.repeat(a[:, :, None], 3, axis=2)
    ctx.upload(0, a)
    xy0, r0 = ctx.fast(0, 20, True)
    ctx.build_pyramid(0, (21, 21), 3)
    l2 = ctx.download_level(0, 2)
    ctx.upload_bgr(1, bgr)
    assert np.array_equal(ctx.download_image(1, a.shape[1], a.shape[0]), a)
    xy1, r1 = ctx.fast(1, 20, True)
    ctx.build_pyramid(1, (21, 21), 3)
    assert np.array_equal(xy0, xy1) and np.array_equal(r0, r1) and len(xy0) > 50
    assert np.array_equal(l2, ctx.download_level(1, 2))


def test_bgr_errors(ctx):
    img = np.zeros((4, 4, 3), np.uint8)
    with pytest.raises(capi.SvoGpuError):
        ctx._ck(ctx.lib.svo_gpu_upload_bgr(ctx.h, 0, capi._ptr(img), 4, 4, 11))          # stride < 3 w
    with pytest.raises(capi.SvoGpuError):
        ctx._ck(ctx.lib.svo_gpu_upload_bgr(ctx.h, 0, None, 4, 4, 12))                    # null image
    with pytest.raises(capi.SvoGpuError):
        ctx._ck(ctx.lib.svo_gpu_upload_bgr(ctx.h, 0, capi._ptr(img), 5000, 4, 15000))    # wider than the context
