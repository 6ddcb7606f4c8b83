"""ctypes binding of include/svo_gpu.h (libsvo_gpu.so) -- the only route from Python to the sm_100a kernels.

This module never falls back to a CPU implementation and never imports anything under oracle/: if the library is missing
or a call fails it raises.  Tests and bench.py go through this binding, i.e. through the C ABI.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# SVO_GPU_LIB selects another build of the same library (kernel tuning variants); never a different implementation
LIB_PATH = os.environ.get("SVO_GPU_LIB") or os.path.join(_HERE, "lib", "libsvo_gpu.so")

OK, E_ARG, E_CAPACITY, E_CUDA, E_UNSUPPORTED, E_STATE = 0, -1, -2, -3, -4, -5
NUM_SLOTS = 5
TC_COUNT, TC_EPS = 1, 2
LK_USE_INITIAL_FLOW, LK_GET_MIN_EIGENVALS = 4, 8
MEM_HOST, MEM_DEVICE = 0, 1

_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int)

# every symbol include/svo_gpu.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "svo_gpu_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "svo_gpu_destroy": (C.c_int, [C.c_void_p]),
    "svo_gpu_last_error": (C.c_char_p, [C.c_void_p]),
    "svo_gpu_stream": (C.c_void_p, [C.c_void_p]),
    "svo_gpu_abi_version": (C.c_int, []),
    "svo_gpu_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "svo_gpu_host_free": (C.c_int, [C.c_void_p]),
    "svo_gpu_upload": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_size_t]),
    "svo_gpu_upload_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t,
                                       C.c_size_t, C.c_int]),
    "svo_gpu_upload_bgr": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_size_t]),
    "svo_gpu_upload_bgr_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t,
                                           C.c_size_t, C.c_int]),
    "svo_gpu_download_image": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t]),
    "svo_gpu_build_pyramid": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, _i32p]),
    "svo_gpu_pyramid_levels": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "svo_gpu_pyramid_margin": (C.c_int, [C.c_int, C.c_int]),
    "svo_gpu_slot_pyramid": (C.c_int, [C.c_void_p, C.c_int, _i32p, _i32p]),
    "svo_gpu_slot_generation": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_ulonglong)]),
    "svo_gpu_download_level": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, _i32p, _i32p]),
    "svo_gpu_download_level_padded": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, _i32p, _i32p,
                                                _i32p]),
    "svo_gpu_fast": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p,
                               C.c_int, _i32p]),
    "svo_gpu_fast_masked_by_points": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_float,
                                                C.c_void_p, C.c_void_p, C.c_int, _i32p]),
    "svo_gpu_bucket": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                 C.c_int, _i32p]),
    "svo_gpu_bucket_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                        C.c_void_p, C.c_int, _i32p, C.c_void_p, C.c_char_p, C.c_size_t]),
    "svo_gpu_klt": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                              C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_double]),
    "svo_gpu_last_klt_iterations": (C.c_longlong, [C.c_void_p]),
    "svo_gpu_filter_tracks": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_int, _i32p]),
    "svo_gpu_track_frames": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_double, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_int, _i32p]),
    "svo_gpu_keyframe": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_int, _i32p, _i32p]),
    "svo_gpu_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_size_t, C.c_size_t, C.c_int, C.c_void_p,
                               C.c_void_p]),
    "svo_gpu_step_pipelined": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_size_t, C.c_size_t, C.c_int,
                                         C.c_void_p, C.c_void_p]),
    "svo_gpu_step_submit": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_size_t, C.c_size_t, C.c_int,
                                      C.c_void_p, C.c_void_p, _i32p]),
    "svo_gpu_step_wait": (C.c_int, [C.c_void_p, C.c_int]),
    "svo_gpu_step_reset": (C.c_int, [C.c_void_p]),
    "svo_gpu_launch_count": (C.c_longlong, [C.c_void_p]),
    "svo_gpu_step_profile": (C.c_int, [C.c_void_p, C.c_int, _f32p]),
}


class StepParams(C.Structure):
    """struct svo_gpu_step_params"""
    _fields_ = [("w", C.c_int), ("h", C.c_int), ("fast_threshold", C.c_int), ("fast_nms", C.c_int),
                ("mask_half", C.c_float), ("bucket_size", C.c_int), ("per_bucket", C.c_int), ("max_tracks", C.c_int),
                ("win_w", C.c_int), ("win_h", C.c_int), ("max_level", C.c_int), ("max_count", C.c_int),
                ("epsilon", C.c_double), ("criteria_type", C.c_int), ("flags", C.c_int),
                ("min_eig_threshold", C.c_double), ("carry_tracks", C.c_int), ("features_to_track", C.c_int)]


class KeyframeParams(C.Structure):
    """struct svo_gpu_keyframe_params"""
    _fields_ = [("fast_threshold", C.c_int), ("fast_nms", C.c_int), ("mask_half", C.c_float), ("bucket_size", C.c_int),
                ("per_bucket", C.c_int), ("win_w", C.c_int), ("win_h", C.c_int), ("max_level", C.c_int),
                ("max_count", C.c_int), ("epsilon", C.c_double), ("criteria_type", C.c_int), ("flags", C.c_int),
                ("min_eig_threshold", C.c_double), ("y_threshold", C.c_float)]


class StepOutputs(C.Structure):
    """struct svo_gpu_step_outputs"""
    _fields_ = [("n_tracked", C.c_void_p), ("next_xy", C.c_void_p), ("status", C.c_void_p), ("err", C.c_void_p),
                ("track_capacity", C.c_int), ("n_keypoints", C.c_void_p), ("kp_xy", C.c_void_p),
                ("kp_response", C.c_void_p), ("kp_capacity", C.c_int), ("n_bucketed", C.c_void_p),
                ("bucket_xy", C.c_void_p), ("bucket_capacity", C.c_int), ("keyframe", C.c_void_p),
                ("n_survived", C.c_void_p), ("survivor_index", C.c_void_p)]


class SvoGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("svo_gpu error %d: %s" % (code, msg))
        self.code = code


_lib = None


def load():
    """Loads libsvo_gpu.so and declares every symbol of the header.  Raises if the library or a symbol is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(make -C svo_b200/csrc).  There is no CPU fallback." % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)   # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def bucket_static(xy_in_ptr, ages_in_ptr, n_in, img_w, img_h, bucket_size, per_bucket, xy_out_ptr, ages_out_ptr, capacity,
                  stream=0):
    """svo_gpu_bucket_static: bucketing on caller-owned DEVICE buffers (raw addresses) with the tables in the library's static
    device array -- no context and no allocation.  Returns the number of bucketed features written to the output buffers."""
    m = C.c_int(0)
    msg = C.create_string_buffer(512)
    rc = load().svo_gpu_bucket_static(xy_in_ptr, ages_in_ptr, n_in, img_w, img_h, bucket_size, per_bucket, xy_out_ptr,
                                      ages_out_ptr, capacity, C.byref(m), stream, msg, len(msg))
    if rc != OK:
        e = SvoGpuError(rc, msg.value.decode("utf-8", "replace"))
        e.required = m.value
        raise e
    return m.value


class Context:
    """RAII wrapper of svo_gpu_ctx."""

    def __init__(self, device=0, max_w=1280, max_h=512, max_levels=4, max_points=32768, max_batch=1):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.svo_gpu_create(C.byref(h), device, max_w, max_h, max_levels, max_points, max_batch)
        if rc != 0:
            raise SvoGpuError(rc, self.lib.svo_gpu_last_error(None).decode())
        self.h = h
        self.max_points = max_points
        self.max_batch = max_batch

    def close(self):
        if getattr(self, "h", None):
            self.lib.svo_gpu_destroy(self.h)
            self.h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc != 0:
            raise SvoGpuError(rc, self.lib.svo_gpu_last_error(self.h).decode())

    @property
    def stream(self):
        return self.lib.svo_gpu_stream(self.h)

    @property
    def launch_count(self):
        return self.lib.svo_gpu_launch_count(self.h)

    # ------------------------------------------------------------------ images
    def upload(self, slot, img):
        img = np.asarray(img)
        if img.dtype != np.uint8 or img.ndim != 2 or img.strides[1] != 1:
            raise ValueError("image must be a 2-D uint8 array with contiguous rows (CV_8UC1)")
        self._ck(self.lib.svo_gpu_upload(self.h, slot, _ptr(img), img.shape[1], img.shape[0], img.strides[0]))

    def upload_batch(self, slot, imgs):
        imgs = np.ascontiguousarray(imgs, np.uint8)
        assert imgs.ndim == 3
        b, h, w = imgs.shape
        self._ck(self.lib.svo_gpu_upload_batch(self.h, slot, _ptr(imgs), MEM_HOST, w, h, w, w * h, b))

    def upload_batch_device(self, slot, dev_ptr, w, h, stride, image_stride, batch):
        self._ck(self.lib.svo_gpu_upload_batch(self.h, slot, C.c_void_p(dev_ptr), MEM_DEVICE, w, h, stride,
                                               image_stride, batch))

    def upload_bgr(self, slot, img):
        """(h, w, 3) uint8 BGR image (rows may be strided) -> gray in the slot, cv::cvtColor(COLOR_BGR2GRAY) on the device."""
        img = np.asarray(img)
        if img.dtype != np.uint8 or img.ndim != 3 or img.shape[2] != 3 or img.strides[2] != 1 or img.strides[1] != 3:
            raise ValueError("image must be an (h, w, 3) uint8 array with interleaved channels (CV_8UC3)")
        self._ck(self.lib.svo_gpu_upload_bgr(self.h, slot, _ptr(img), img.shape[1], img.shape[0], img.strides[0]))

    def upload_bgr_batch(self, slot, imgs):
        imgs = np.ascontiguousarray(imgs, np.uint8)
        assert imgs.ndim == 4 and imgs.shape[3] == 3
        b, h, w, _ = imgs.shape
        self._ck(self.lib.svo_gpu_upload_bgr_batch(self.h, slot, _ptr(imgs), MEM_HOST, w, h, 3 * w, 3 * w * h, b))

    def upload_bgr_batch_device(self, slot, dev_ptr, w, h, stride, image_stride, batch):
        self._ck(self.lib.svo_gpu_upload_bgr_batch(self.h, slot, C.c_void_p(dev_ptr), MEM_DEVICE, w, h, stride,
                                                   image_stride, batch))

    def download_image(self, slot, w, h, lane=0):
        out = np.empty((h, w), np.uint8)
        self._ck(self.lib.svo_gpu_download_image(self.h, slot, lane, _ptr(out), out.strides[0]))
        return out

    def build_pyramid(self, slot, win=(21, 21), max_level=3):
        lv = C.c_int(0)
        self._ck(self.lib.svo_gpu_build_pyramid(self.h, slot, win[0], win[1], max_level, C.byref(lv)))
        return lv.value

    def download_level(self, slot, level, lane=0):
        w, h = C.c_int(0), C.c_int(0)
        rc = self.lib.svo_gpu_download_level(self.h, slot, lane, level, None, 0, C.byref(w), C.byref(h))
        if rc not in (0, E_CAPACITY):
            self._ck(rc)
        out = np.empty((h.value, w.value), np.uint8)
        self._ck(self.lib.svo_gpu_download_level(self.h, slot, lane, level, _ptr(out), out.size, C.byref(w), C.byref(h)))
        return out

    def download_level_padded(self, slot, level, lane=0):
        """Level with its REFLECT_101 margin -> (array (h + 2m, w + 2m), m)."""
        w, h, m = C.c_int(0), C.c_int(0), C.c_int(0)
        rc = self.lib.svo_gpu_download_level_padded(self.h, slot, lane, level, None, 0, C.byref(w), C.byref(h), C.byref(m))
        if rc not in (0, E_CAPACITY):
            self._ck(rc)
        out = np.empty((h.value + 2 * m.value, w.value + 2 * m.value), np.uint8)
        self._ck(self.lib.svo_gpu_download_level_padded(self.h, slot, lane, level, _ptr(out), out.size, C.byref(w),
                                                        C.byref(h), C.byref(m)))
        return out, m.value

    # ------------------------------------------------------------------ FAST
    def fast(self, slot, threshold=20, nms=True, mask=None, capacity=None):
        cap = self.max_points if capacity is None else capacity
        xy = np.empty((max(cap, 1), 2), np.float32)
        resp = np.empty(max(cap, 1), np.float32)
        k = C.c_int(0)
        if mask is not None:
            mask = np.asarray(mask)
            if mask.dtype != np.uint8 or mask.ndim != 2 or mask.strides[1] != 1:
                raise ValueError("mask must be CV_8UC1")
        self._ck(self.lib.svo_gpu_fast(self.h, slot, int(threshold), int(bool(nms)), _ptr(mask),
                                       mask.strides[0] if mask is not None else 0, _ptr(xy), _ptr(resp), cap, C.byref(k)))
        return xy[:k.value].copy(), resp[:k.value].copy()

    def fast_masked_by_points(self, slot, pts, half=10.0, threshold=20, nms=True, capacity=None):
        cap = self.max_points if capacity is None else capacity
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
        xy = np.empty((max(cap, 1), 2), np.float32)
        resp = np.empty(max(cap, 1), np.float32)
        k = C.c_int(0)
        self._ck(self.lib.svo_gpu_fast_masked_by_points(self.h, slot, int(threshold), int(bool(nms)), _ptr(pts), len(pts),
                                                        float(half), _ptr(xy), _ptr(resp), cap, C.byref(k)))
        return xy[:k.value].copy(), resp[:k.value].copy()

    # ------------------------------------------------------------------ bucketing
    def bucket(self, xy, ages, img_w, img_h, bucket_size, per_bucket, capacity=None):
        xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
        ages = np.ascontiguousarray(ages, np.int32).reshape(-1)
        n = len(xy)
        assert len(ages) == n
        if capacity is None:
            capacity = max(n, (img_h // max(bucket_size, 1) + 1) * (img_w // max(bucket_size, 1) + 1) * max(per_bucket, 1))
        bx = np.zeros((max(capacity, 1), 2), np.float32)
        ba = np.zeros(max(capacity, 1), np.int32)
        bx[:n] = xy
        ba[:n] = ages
        m = C.c_int(0)
        self._ck(self.lib.svo_gpu_bucket(self.h, _ptr(bx), _ptr(ba), n, img_w, img_h, bucket_size, per_bucket, capacity,
                                         C.byref(m)))
        return bx[:m.value].copy(), ba[:m.value].copy()

    # ------------------------------------------------------------------ LK
    def klt(self, slot_prev, slot_next, prev_xy, win=(21, 21), max_level=3, max_count=30, epsilon=0.01,
            criteria_type=TC_COUNT | TC_EPS, flags=0, min_eig_threshold=1e-4):
        pxy = np.ascontiguousarray(prev_xy, np.float32).reshape(-1, 2)
        n = len(pxy)
        nxy = np.zeros((max(n, 1), 2), np.float32)
        st = np.zeros(max(n, 1), np.uint8)
        er = np.zeros(max(n, 1), np.float32)
        self._ck(self.lib.svo_gpu_klt(self.h, slot_prev, slot_next, _ptr(pxy), _ptr(nxy), _ptr(st), _ptr(er), n, win[0],
                                      win[1], max_level, max_count, float(epsilon), criteria_type, flags,
                                      float(min_eig_threshold)))
        return nxy[:n], st[:n], er[:n]

    # ------------------------------------------------------------------ filters and fused call chains
    def filter_tracks(self, a_xy, b_xy, status, y_threshold=-1.0):
        """src/tracking.cpp:168-175 (y_threshold < 0) / :109-115 -> (index, a, b) of the survivors, input order kept."""
        a = np.ascontiguousarray(a_xy, np.float32).reshape(-1, 2)
        b = np.ascontiguousarray(b_xy, np.float32).reshape(-1, 2)
        st = np.ascontiguousarray(status, np.uint8).reshape(-1)
        n = len(a)
        assert len(b) == n and len(st) == n
        idx = np.zeros(max(n, 1), np.int32)
        ao = np.zeros((max(n, 1), 2), np.float32)
        bo = np.zeros((max(n, 1), 2), np.float32)
        m = C.c_int(0)
        self._ck(self.lib.svo_gpu_filter_tracks(self.h, _ptr(a), _ptr(b), _ptr(st), n, float(y_threshold), _ptr(idx),
                                                _ptr(ao), _ptr(bo), n, C.byref(m)))
        return idx[:m.value].copy(), ao[:m.value].copy(), bo[:m.value].copy()

    def track_frames(self, slot_prev, slot_next, prev_xy, win=(21, 21), max_level=3, max_count=50, epsilon=0.001,
                     criteria_type=TC_COUNT | TC_EPS, flags=LK_GET_MIN_EIGENVALS, min_eig_threshold=1e-4):
        """Tracking::trackFrames (src/tracking.cpp:154-179): LK + status filter -> (index, curr_xy, err) of the survivors."""
        pxy = np.ascontiguousarray(prev_xy, np.float32).reshape(-1, 2)
        n = len(pxy)
        idx = np.zeros(max(n, 1), np.int32)
        cur = np.zeros((max(n, 1), 2), np.float32)
        er = np.zeros(max(n, 1), np.float32)
        m = C.c_int(0)
        self._ck(self.lib.svo_gpu_track_frames(self.h, slot_prev, slot_next, _ptr(pxy), n, win[0], win[1], max_level,
                                               max_count, float(epsilon), criteria_type, flags, float(min_eig_threshold),
                                               _ptr(idx), _ptr(cur), _ptr(er), n, C.byref(m)))
        return idx[:m.value].copy(), cur[:m.value].copy(), er[:m.value].copy()

    def keyframe(self, slot_left, slot_right, prev_xy, fast_threshold=20, fast_nms=True, mask_half=10.0, bucket_size=0,
                 per_bucket=0, win=(11, 11), max_level=3, max_count=30, epsilon=0.001, criteria_type=TC_COUNT | TC_EPS,
                 flags=0, min_eig_threshold=1e-4, y_threshold=40.0, capacity=None):
        """Tracking::extractFeatures + findLeftFeaturesInRight (src/tracking.cpp:74-122) -> (left, right, n_detected)."""
        pxy = np.ascontiguousarray(prev_xy, np.float32).reshape(-1, 2)
        cap = self.max_points if capacity is None else capacity
        P = KeyframeParams(int(fast_threshold), int(bool(fast_nms)), float(mask_half), int(bucket_size), int(per_bucket),
                           win[0], win[1], max_level, max_count, float(epsilon), criteria_type, flags,
                           float(min_eig_threshold), float(y_threshold))
        lo = np.zeros((max(cap, 1), 2), np.float32)
        ro = np.zeros((max(cap, 1), 2), np.float32)
        nd, m = C.c_int(0), C.c_int(0)
        self._ck(self.lib.svo_gpu_keyframe(self.h, slot_left, slot_right, _ptr(pxy) if len(pxy) else None, len(pxy),
                                           C.byref(P), _ptr(lo), _ptr(ro), cap, C.byref(nd), C.byref(m)))
        return lo[:m.value].copy(), ro[:m.value].copy(), nd.value

    @property
    def last_klt_iterations(self):
        return self.lib.svo_gpu_last_klt_iterations(self.h)

    # ------------------------------------------------------------------ per-frame step
    def step(self, images_ptr, mem_kind, stride, image_stride, batch, params, outputs=None):
        self._ck(self.lib.svo_gpu_step(self.h, C.c_void_p(images_ptr), mem_kind, stride, image_stride, batch,
                                       C.byref(params), C.byref(outputs) if outputs is not None else None))

    def step_pipelined(self, images_ptr, next_ptr, mem_kind, stride, image_stride, batch, params, outputs=None):
        """next_ptr: host address of the FOLLOWING step's frames (copied during this step) or None."""
        self._ck(self.lib.svo_gpu_step_pipelined(self.h, C.c_void_p(images_ptr), C.c_void_p(next_ptr) if next_ptr else None,
                                                 mem_kind, stride, image_stride, batch, C.byref(params),
                                                 C.byref(outputs) if outputs is not None else None))

    def step_submit(self, images_ptr, next_ptr, mem_kind, stride, image_stride, batch, params, outputs=None):
        """Enqueues the step and returns its ticket; up to STEP_DEPTH may be outstanding (see svo_gpu.h)."""
        t = C.c_int(-1)
        self._ck(self.lib.svo_gpu_step_submit(self.h, C.c_void_p(images_ptr), C.c_void_p(next_ptr) if next_ptr else None,
                                              mem_kind, stride, image_stride, batch, C.byref(params),
                                              C.byref(outputs) if outputs is not None else None, C.byref(t)))
        return t.value

    def step_wait(self, ticket):
        self._ck(self.lib.svo_gpu_step_wait(self.h, int(ticket)))

    def step_reset(self):
        self._ck(self.lib.svo_gpu_step_reset(self.h))

    def step_profile(self, enable=True):
        ms = (C.c_float * 4)()
        self._ck(self.lib.svo_gpu_step_profile(self.h, int(enable), ms))
        return list(ms)
