"""ctypes loader for the CPU oracle.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs may import this module,
and only as the checker or the timed CPU baseline -- never as a fallback for the CUDA path (svo_b200 never imports it).

Three layers, strongest first:
  * ``cv2_*``      -- the OpenCV the reference itself calls (src/tracking.cpp:82, :101-105, :160-165), through the
                      cv2 wheel of this image, when importable.  This is "the reference's own CPU implementation".
  * ``ref_bucket`` -- the reference's own src/bucket.cpp compiled unmodified (oracle/_ref/libref_bucket.so).
  * ``orc_*``      -- oracle/svo_oracle.c, the plain-C restatement (pinned against the two above by
                      tests/test_oracle_pin.py and the fixtures under tests/golden/).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libsvo_oracle.so")
_REF = os.path.join(_HERE, "_ref", "libref_bucket.so")

LK_GET_MIN_EIGENVALS = 8
TC_COUNT, TC_EPS = 1, 2

_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int)


def build(force=False):
    """Compile the C restatement (and oracle/_ref when /root/reference is present)."""
    stale = os.path.exists(_LIB) and os.path.getmtime(os.path.join(_HERE, "svo_oracle.c")) > os.path.getmtime(_LIB)
    if force or stale or not os.path.exists(_LIB) or (os.path.isdir("/root/reference/src") and not os.path.exists(_REF)):
        import sys
        subprocess.check_call(["make", "-C", _HERE, "-s"], stdout=sys.stderr)   # bench.py's stdout carries one JSON line only


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB)
        _lib.orc_fast.restype = C.c_int
        _lib.orc_fast.argtypes = [_u8p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, _u8p, C.c_size_t,
                                  _f32p, _f32p, C.c_int]
        _lib.orc_build_mask.restype = None
        _lib.orc_build_mask.argtypes = [_u8p, C.c_int, C.c_int, C.c_size_t, _f32p, C.c_int, C.c_float]
        _lib.orc_bucket.restype = C.c_int
        _lib.orc_bucket.argtypes = [_f32p, _i32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _i32p]
        _lib.orc_pyrdown.restype = None
        _lib.orc_bgr2gray.argtypes = [_u8p, C.c_int, C.c_int, C.c_size_t, _u8p, C.c_size_t]
        _lib.orc_bgr2gray.restype = None
        _lib.orc_pyrdown.argtypes = [_u8p, C.c_int, C.c_int, C.c_size_t, _u8p, C.c_size_t]
        _lib.orc_scharr.restype = None
        _lib.orc_scharr.argtypes = [_u8p, C.c_int, C.c_int, C.c_size_t, C.POINTER(C.c_int16)]
        _lib.orc_pyramid_levels.restype = C.c_int
        _lib.orc_pyramid_levels.argtypes = [C.c_int] * 5
        _lib.orc_klt_ex.restype = C.c_int
        _lib.orc_klt_ex.argtypes = [_u8p, _u8p, C.c_int, C.c_int, C.c_size_t, C.c_size_t, _f32p, _f32p, _u8p, _f32p,
                                    C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int,
                                    C.c_double, C.c_int, C.POINTER(C.c_longlong)]
    return _lib


def have_ref_bucket():
    build()
    return os.path.exists(_REF)


def ref_lib():
    global _ref
    if _ref is None:
        build()
        _ref = C.CDLL(_REF)
        _ref.ref_bucketing.restype = C.c_int
        _ref.ref_bucketing.argtypes = [_f32p, _i32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _i32p]
        _ref.ref_append_then_size.restype = C.c_int
        _ref.ref_append_then_size.argtypes = [_f32p, C.c_int]
    return _ref


def _img(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    assert a.ndim == 2
    return a


def _p(a, t):
    return a.ctypes.data_as(t)


# ------------------------------------------------------------------------------------------------ C restatement
def orc_fast(img, threshold=20, nms=True, mask=None):
    """-> (xy float32 [K,2], response float32 [K]) in raster order."""
    img = _img(img)
    h, w = img.shape
    cap = max(1, w * h)
    xy = np.empty((cap, 2), np.float32)
    resp = np.empty(cap, np.float32)
    m = _img(mask) if mask is not None else None
    k = lib().orc_fast(_p(img, _u8p), w, h, img.strides[0], int(threshold), int(bool(nms)),
                       _p(m, _u8p) if m is not None else None, m.strides[0] if m is not None else 0,
                       _p(xy, _f32p), _p(resp, _f32p), cap)
    return xy[:k].copy(), resp[:k].copy()


def orc_build_mask(w, h, xy, half=10.0):
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    mask = np.empty((h, w), np.uint8)
    lib().orc_build_mask(_p(mask, _u8p), w, h, mask.strides[0], _p(xy, _f32p), len(xy), float(half))
    return mask


def _bucket(fn, xy, ages, img_w, img_h, bucket_size, per_bucket):
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    n = len(xy)
    ages = np.ascontiguousarray(ages, np.int32).reshape(-1)
    assert len(ages) == n
    hc, wc = img_h // bucket_size, img_w // bucket_size
    cap = max(n, (hc + 1) * (wc + 1) * max(per_bucket, 1)) + 1
    bx = np.zeros((cap, 2), np.float32)
    ba = np.zeros(cap, np.int32)
    bx[:n] = xy
    ba[:n] = ages
    nout = C.c_int(0)
    rc = fn(_p(bx, _f32p), _p(ba, _i32p), n, img_w, img_h, bucket_size, per_bucket, cap, C.byref(nout))
    if rc != 0:
        raise ValueError("bucketing failed rc=%d" % rc)
    return bx[:nout.value].copy(), ba[:nout.value].copy()


def orc_bucket(xy, ages, img_w, img_h, bucket_size, per_bucket):
    return _bucket(lib().orc_bucket, xy, ages, img_w, img_h, bucket_size, per_bucket)


def ref_bucket(xy, ages, img_w, img_h, bucket_size, per_bucket):
    """The reference's own FeatureSet::bucketingFeatures (src/bucket.cpp:24-68), compiled unmodified."""
    return _bucket(ref_lib().ref_bucketing, xy, ages, img_w, img_h, bucket_size, per_bucket)


def orc_bgr2gray(bgr):
    """cv::cvtColor(bgr, COLOR_BGR2GRAY) of an (h, w, 3) uint8 image (include/async_image_loader.h:68-69)."""
    bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
    h, w, _ = bgr.shape
    out = np.empty((h, w), np.uint8)
    lib().orc_bgr2gray(_p(bgr, _u8p), w, h, bgr.strides[0], _p(out, _u8p), out.strides[0])
    return out


def orc_pyrdown(img):
    img = _img(img)
    h, w = img.shape
    out = np.empty(((h + 1) // 2, (w + 1) // 2), np.uint8)
    lib().orc_pyrdown(_p(img, _u8p), w, h, img.strides[0], _p(out, _u8p), out.strides[0])
    return out


def orc_pyramid(img, win, max_level):
    """Image levels 0..L of buildOpticalFlowPyramid (unpadded), L = effective max level."""
    img = _img(img)
    h, w = img.shape
    L = lib().orc_pyramid_levels(w, h, win[0], win[1], max_level)
    out = [img]
    for _ in range(L):
        out.append(orc_pyrdown(out[-1]))
    return out


def orc_scharr(img):
    img = _img(img)
    h, w = img.shape
    d = np.empty((h, w, 2), np.int16)
    lib().orc_scharr(_p(img, _u8p), w, h, img.strides[0], _p(d, C.POINTER(C.c_int16)))
    return d


def orc_klt(prev, nxt, prev_xy, win=(21, 21), max_level=3, max_count=30, epsilon=0.01,
            criteria_type=TC_COUNT | TC_EPS, flags=0, min_eig_threshold=1e-4, accum_mode=0, return_iters=False):
    """-> (next_xy [n,2] f32, status [n] u8, err [n] f32)  (cv::calcOpticalFlowPyrLK semantics)."""
    prev, nxt = _img(prev), _img(nxt)
    assert prev.shape == nxt.shape
    h, w = prev.shape
    pxy = np.ascontiguousarray(prev_xy, np.float32).reshape(-1, 2)
    n = len(pxy)
    nxy = np.zeros((max(n, 1), 2), np.float32)
    st = np.zeros(max(n, 1), np.uint8)
    er = np.zeros(max(n, 1), np.float32)
    it = C.c_longlong(0)
    rc = lib().orc_klt_ex(_p(prev, _u8p), _p(nxt, _u8p), w, h, prev.strides[0], nxt.strides[0], _p(pxy, _f32p),
                          _p(nxy, _f32p), _p(st, _u8p), _p(er, _f32p), n, win[0], win[1], max_level, max_count,
                          float(epsilon), criteria_type, flags, float(min_eig_threshold), accum_mode, C.byref(it))
    if rc != 0:
        raise ValueError("orc_klt rc=%d" % rc)
    if return_iters:
        return nxy[:n], st[:n], er[:n], it.value
    return nxy[:n], st[:n], er[:n]


# ------------------------------------------------------------------------------------------------ OpenCV itself
def have_cv2():
    try:
        import cv2  # noqa: F401
        return True
    except Exception:
        return False


def cv2_fast(img, threshold=20, nms=True, mask=None, with_response=True):
    """cv::FastFeatureDetector::create(threshold, nms)->detect(img, kps, mask) + cv::KeyPoint::convert
    (src/tracking.cpp:54-57, :82, :86).  The responses are read keypoint by keypoint in Python: parity tests only (the
    timed CPU arm, oracle/cpu_arm.py, never asks for them)."""
    import cv2
    det = cv2.FastFeatureDetector_create(int(threshold), bool(nms))
    kps = det.detect(_img(img), _img(mask) if mask is not None else None)
    xy = cv2.KeyPoint_convert(kps).reshape(-1, 2).astype(np.float32) if len(kps) else np.zeros((0, 2), np.float32)
    resp = np.array([k.response for k in kps], np.float32) if with_response else None
    return xy, resp


def cv2_build_mask(w, h, xy, half=10.0):
    """src/tracking.cpp:76-79."""
    import cv2
    mask = np.full((h, w), 255, np.uint8)
    xy = np.asarray(xy, np.float32).reshape(-1, 2)
    p1 = np.rint(xy - np.float32(half)).astype(np.int64).tolist()   # float32 subtraction, then round-half-even (Point2f -> Point)
    p2 = np.rint(xy + np.float32(half)).astype(np.int64).tolist()
    for a, b in zip(p1, p2):
        cv2.rectangle(mask, tuple(a), tuple(b), 0, cv2.FILLED)
    return mask


def cv2_klt(prev, nxt, prev_xy, win=(21, 21), max_level=3, max_count=30, epsilon=0.01,
            criteria_type=TC_COUNT | TC_EPS, flags=0, min_eig_threshold=1e-4):
    """cv::calcOpticalFlowPyrLK as called at src/tracking.cpp:101-105 / :160-165."""
    import cv2
    pxy = np.ascontiguousarray(prev_xy, np.float32).reshape(-1, 1, 2)
    if len(pxy) == 0:
        return np.zeros((0, 2), np.float32), np.zeros(0, np.uint8), np.zeros(0, np.float32)
    nxy, st, er = cv2.calcOpticalFlowPyrLK(_img(prev), _img(nxt), pxy, None, winSize=tuple(win), maxLevel=max_level,
                                           criteria=(criteria_type, max_count, epsilon), flags=flags,
                                           minEigThreshold=min_eig_threshold)
    return nxy.reshape(-1, 2), st.reshape(-1), er.reshape(-1)
