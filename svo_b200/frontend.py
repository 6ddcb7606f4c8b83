"""Host-side driver of the batched per-frame step (svo_gpu_step) for tests and bench.py.

Mirrors the loop body of Tracking::startStereo (reference src/tracking.cpp:240-257) for `batch` independent streams
on one GPU.  Everything numeric happens in libsvo_gpu.so; this file only owns the host result buffers.
"""
import ctypes as C

import numpy as np

from . import capi

REFERENCE_PARAMS = dict(fast_threshold=20, fast_nms=1, mask_half=10.0, bucket_size=32, per_bucket=5, max_tracks=2000,
                        win_w=21, win_h=21, max_level=3, max_count=50, epsilon=0.001, criteria_type=3,
                        flags=capi.LK_GET_MIN_EIGENVALS, min_eig_threshold=1e-4, carry_tracks=0, features_to_track=70)


def make_params(w, h, **over):
    d = dict(REFERENCE_PARAMS, w=w, h=h)
    d.update(over)
    p = capi.StepParams()
    for k, v in d.items():
        setattr(p, k, v)
    return p, d


def _pinned(shape, dtype):
    """Pinned host array when torch+CUDA are around (plumbing only), else plain numpy."""
    try:
        import torch
        if torch.cuda.is_available():
            t = torch.empty(shape, dtype=getattr(torch, np.dtype(dtype).name), pin_memory=True)
            return t.numpy(), t
    except Exception:
        pass
    return np.empty(shape, dtype), None


STEP_DEPTH = 2   # SVO_GPU_STEP_DEPTH


class _Outputs:
    """One set of (pinned) host result arrays + the svo_gpu_step_outputs view of them."""

    def __init__(self, batch, tc, kp_capacity, bucket_capacity):
        self._keep = []
        self.out = capi.StepOutputs()

        def buf(shape, dtype):
            a, t = _pinned(shape, dtype)
            self._keep.append(t)
            return a

        self.n_tracked = buf((batch,), np.int32)
        self.n_keypoints = buf((batch,), np.int32)
        self.n_bucketed = buf((batch,), np.int32)
        self.keyframe = buf((batch,), np.int32)
        self.n_survived = buf((batch,), np.int32)
        self.survivor_index = buf((batch, max(tc, 1)), np.int32)
        self.next_xy = buf((batch, max(tc, 1), 2), np.float32)
        self.status = buf((batch, max(tc, 1)), np.uint8)
        self.err = buf((batch, max(tc, 1)), np.float32)
        self.bucket_xy = buf((batch, max(bucket_capacity, 1), 2), np.float32)
        o = self.out
        o.n_tracked, o.n_keypoints, o.n_bucketed = (self.n_tracked.ctypes.data, self.n_keypoints.ctypes.data,
                                                    self.n_bucketed.ctypes.data)
        o.next_xy, o.status, o.err = self.next_xy.ctypes.data, self.status.ctypes.data, self.err.ctypes.data
        o.track_capacity = tc
        o.bucket_xy = self.bucket_xy.ctypes.data
        o.bucket_capacity = bucket_capacity
        o.keyframe, o.n_survived, o.survivor_index = (self.keyframe.ctypes.data, self.n_survived.ctypes.data,
                                                      self.survivor_index.ctypes.data)
        self.kp_xy = self.kp_response = None
        if kp_capacity > 0:
            self.kp_xy = buf((batch, kp_capacity, 2), np.float32)
            self.kp_response = buf((batch, kp_capacity), np.float32)
            o.kp_xy, o.kp_response = self.kp_xy.ctypes.data, self.kp_response.ctypes.data
        o.kp_capacity = kp_capacity


class Frontend:
    def __init__(self, ctx, w, h, batch, track_capacity=None, kp_capacity=0, bucket_capacity=4096, **over):
        self.ctx = ctx
        self.batch = batch
        self.params, self.pdict = make_params(w, h, **over)
        tc = self.pdict["max_tracks"] if track_capacity is None else track_capacity
        # one result set per step that may be outstanding (submit / wait); the synchronous calls use set 0
        self._sets = [_Outputs(batch, tc, kp_capacity, bucket_capacity) for _ in range(STEP_DEPTH)]
        self._pending = []
        self._nsub = 0
        self._bind(self._sets[0])
        self.d2h_bytes = int(5 * 4 * batch + batch * tc * (8 + 1 + 4) + batch * bucket_capacity * 8 +
                             batch * kp_capacity * 12 + (batch * tc * 4 if self.pdict["carry_tracks"] else 0))

    def _bind(self, s):
        """The public result attributes show set `s` (the step that completed last)."""
        self.out = s.out
        for name in ("n_tracked", "n_keypoints", "n_bucketed", "next_xy", "status", "err", "bucket_xy", "kp_xy", "kp_response",
                     "keyframe", "n_survived", "survivor_index"):
            setattr(self, name, getattr(s, name))

    # ---- asynchronous form: submit(t + 1) may precede wait(t); results of a step are visible after ITS wait
    def submit(self, ptr, next_ptr, mem_kind, stride, image_stride, with_outputs=True):
        s = self._sets[self._nsub % STEP_DEPTH]
        t = self.ctx.step_submit(ptr, next_ptr, mem_kind, stride, image_stride, self.batch, self.params,
                                 s.out if with_outputs else None)
        self._pending.append((t, s))
        self._nsub += 1
        return t

    def wait(self):
        """Waits for the oldest outstanding step and shows its results."""
        t, s = self._pending.pop(0)
        self.ctx.step_wait(t)
        self._bind(s)

    @property
    def outstanding(self):
        return len(self._pending)

    def reset(self):
        self.ctx.step_reset()
        self._pending = []
        self._bind(self._sets[0])

    def step_host(self, images):
        """images: uint8 [batch, h, w] host array (C-contiguous).  Results land in this object's host arrays."""
        assert images.dtype == np.uint8 and images.ndim == 3 and images.flags.c_contiguous
        b, h, w = images.shape
        self._bind(self._sets[0])
        self.ctx.step(images.ctypes.data, capi.MEM_HOST, w, w * h, b, self.params, self.out)

    def step_host_ptr(self, ptr, stride, image_stride, with_outputs=True):
        self._bind(self._sets[0])
        self.ctx.step(ptr, capi.MEM_HOST, stride, image_stride, self.batch, self.params,
                      self.out if with_outputs else None)

    def step_host_pipelined(self, ptr, next_ptr, stride, image_stride, with_outputs=True):
        """Like step_host_ptr, and starts the copy of the next step's frames (host address next_ptr, or None)."""
        self._bind(self._sets[0])
        self.ctx.step_pipelined(ptr, next_ptr, capi.MEM_HOST, stride, image_stride, self.batch, self.params,
                                self.out if with_outputs else None)

    def step_device(self, dev_ptr, stride, image_stride, with_outputs=False):
        """dev_ptr: device address of `batch` images (row pitch `stride`, image pitch `image_stride`)."""
        self._bind(self._sets[0])
        self.ctx.step(dev_ptr, capi.MEM_DEVICE, stride, image_stride, self.batch, self.params,
                      self.out if with_outputs else None)
