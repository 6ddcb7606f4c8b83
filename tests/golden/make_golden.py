#!/usr/bin/env python
"""Generates the golden fixtures of tests/golden/ from the reference's own CPU code, in the build container:

  * OpenCV (cv2 wheel of this image; the reference calls the same C++ code at src/tracking.cpp:82, :101-105, :160-165,
    mask at :76-79) for FAST, the mask, buildOpticalFlowPyramid / Scharr and calcOpticalFlowPyrLK;
  * the reference's src/bucket.cpp compiled unmodified (oracle/_ref/libref_bucket.so, built from /root/reference by
    oracle/Makefile) for FeatureSet::bucketingFeatures.

The fixtures carry their INPUTS too (images, points), so neither cv2 nor /root/reference is needed to check against them:
the GPU box and the CPU test-suite only read the .npz files.  Re-run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import cv2  # noqa: E402

from oracle import oracle as O  # noqa: E402
from svo_b200 import synth  # noqa: E402


def small_pair(w, h, seed):
    c = synth.make_canvas(w, h, seed)
    return synth.render_frame(c, w, h, 0), synth.render_frame(c, w, h, 1)


def golden_fast():
    out = {}
    a, _ = small_pair(200, 150, 21)
    imgs = {"synth": a, "noise": synth.noise_image(97, 61, 22), "blocks": synth.block_image(120, 80, 23),
            "tiny7": synth.noise_image(7, 7, 24), "thin": synth.noise_image(40, 6, 25)}
    cases = []
    for name, img in imgs.items():
        out["img_" + name] = img
        for thr in (20, 7):
            for nms in (1, 0):
                xy, resp = O.cv2_fast(img, thr, bool(nms))
                key = "%s_t%d_n%d" % (name, thr, nms)
                out["xy_" + key] = xy.astype(np.int16)
                out["resp_" + key] = resp.astype(np.uint8)
                cases.append(key)
    # mask built the reference's way (src/tracking.cpp:76-79) + masked detect
    h, w = a.shape
    pts = np.random.default_rng(5).uniform([-5, -5], [w + 5, h + 5], size=(40, 2)).astype(np.float32)
    pts[:6] = np.round(pts[:6]) + 0.5
    mask = O.cv2_build_mask(w, h, pts, 10.0)
    xy, resp = O.cv2_fast(a, 20, True, mask)
    out.update(mask_pts=pts, mask=mask, xy_masked=xy.astype(np.int16), resp_masked=resp.astype(np.uint8))
    out["cases"] = np.array(cases)
    np.savez_compressed(os.path.join(HERE, "fast.npz"), **out)
    return len(cases)


def golden_pyramid():
    img = synth.noise_image(157, 101, 31)
    out = {"img": img}
    _, pyr = cv2.buildOpticalFlowPyramid(img, (21, 21), 3, None, False)
    for l, p in enumerate(pyr):
        out["level%d" % l] = np.ascontiguousarray(p)
    out["nlevels"] = np.array(len(pyr))
    # a case where construction stops early (level would be <= winSize)
    img2 = synth.noise_image(90, 70, 32)
    ml, pyr2 = cv2.buildOpticalFlowPyramid(img2, (21, 21), 4, None, False)
    out["img2"] = img2
    out["maxlevel2"] = np.array(ml)
    sc = np.stack([cv2.Scharr(img, cv2.CV_16S, 1, 0), cv2.Scharr(img, cv2.CV_16S, 0, 1)], axis=-1)
    out["scharr"] = sc
    np.savez_compressed(os.path.join(HERE, "pyramid.npz"), **out)


def golden_bgr():
    """cv::cvtColor(BGR2GRAY) as the reference's loader applies it (include/async_image_loader.h:68-69)."""
    rng = np.random.default_rng(51)
    rnd = rng.integers(0, 256, size=(37, 61, 3), dtype=np.uint8)                     # odd sizes: ragged row tails
    vals = np.array([0, 1, 2, 127, 128, 129, 254, 255], np.uint8)
    B, G, R = np.meshgrid(np.arange(256, dtype=np.uint8), vals, vals, indexing="ij")
    extreme = np.ascontiguousarray(np.stack([B, G, R], -1).reshape(256, 64, 3))       # every B x extreme G, R
    out = {"rnd": rnd, "rnd_gray": cv2.cvtColor(rnd, cv2.COLOR_BGR2GRAY),
           "extreme": extreme, "extreme_gray": cv2.cvtColor(extreme, cv2.COLOR_BGR2GRAY)}
    np.savez_compressed(os.path.join(HERE, "bgr.npz"), **out)


def golden_klt():
    a, b = small_pair(200, 150, 41)
    xy, _ = O.cv2_fast(a, 20, True)
    rng = np.random.default_rng(7)
    sel = np.sort(rng.permutation(len(xy))[:150])
    pts = xy[sel] + rng.uniform(-0.5, 0.5, (len(sel), 2)).astype(np.float32)
    h, w = a.shape
    adv = np.array([[0, 0], [w - 1, h - 1], [w - 0.5, 10], [5, h - 2], [-30, -30], [w + 25, h + 25], [-12.3, 50.2],
                    [w / 2, -9.5], [3.25, 3.75], [w - 3.5, h - 3.5]], np.float32)
    pts = np.concatenate([pts, adv]).astype(np.float32)
    out = {"prev": a, "next": b, "pts": pts}
    modes = {"temporal": dict(win=(21, 21), max_level=3, max_count=50, epsilon=0.001, flags=8),   # tracking.cpp:160-165
             "stereo": dict(win=(11, 11), max_level=3, max_count=30, epsilon=0.001, flags=0)}     # tracking.cpp:101-105
    for name, m in modes.items():
        nxy, st, er = O.cv2_klt(a, b, pts, **m)
        out["next_" + name], out["status_" + name], out["err_" + name] = nxy, st, er
    np.savez_compressed(os.path.join(HERE, "klt.npz"), **out)


def golden_bucket():
    rng = np.random.default_rng(0)
    out = {}
    cases = {
        "toy": (np.array([[1, 1], [2, 2], [3, 3], [4, 4], [5, 5]], np.float32), np.array([5, 1, 9, 0, 7], np.int32), 100, 100, 50, 2),
        "uniform5000": (rng.uniform([0, 0], [1241, 376], (5000, 2)).astype(np.float32), rng.integers(0, 50, 5000).astype(np.int32), 1241, 376, 50, 4),
        "k8": (rng.uniform([0, 0], [752, 480], (3000, 2)).astype(np.float32), rng.integers(0, 9, 3000).astype(np.int32), 752, 480, 24, 8),
        "dense": (rng.uniform([0, 0], [200, 100], (4000, 2)).astype(np.float32), rng.integers(0, 100, 4000).astype(np.int32), 200, 100, 10, 3),
    }
    for name, (xy, ages, w, h, bs, k) in cases.items():
        rx, ra = O.ref_bucket(xy, ages, w, h, bs, k)
        out["in_xy_" + name], out["in_age_" + name] = xy, ages
        out["cfg_" + name] = np.array([w, h, bs, k])
        out["out_xy_" + name], out["out_age_" + name] = rx, ra
    out["cases"] = np.array(list(cases))
    np.savez_compressed(os.path.join(HERE, "bucket.npz"), **out)


if __name__ == "__main__":
    O.build()
    assert O.have_ref_bucket(), "oracle/_ref/libref_bucket.so missing (needs /root/reference)"
    n = golden_fast()
    golden_pyramid()
    golden_bgr()
    golden_klt()
    golden_bucket()
    with open(os.path.join(HERE, "VERSIONS.txt"), "w") as f:
        f.write("cv2 %s\nnumpy %s\nreference bucket.cpp: /root/reference/src/bucket.cpp (compiled unmodified)\n" %
                (cv2.__version__, np.__version__))
    print("golden fixtures written:", n, "FAST cases;", sorted(os.listdir(HERE)))
