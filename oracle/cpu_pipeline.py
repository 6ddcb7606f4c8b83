"""CPU per-frame front-end step.  TEST INFRASTRUCTURE ONLY (see oracle/oracle.py): the checker for svo_gpu_step in
tests/, and the timed CPU arm of bench.py (`cpu_baseline` and `--impl reference`).  Never imported by svo_b200.

One step = what Tracking::startStereo does per key-frame (reference src/tracking.cpp:240-257):
  trackFrames(prev, curr)  -> cv::calcOpticalFlowPyrLK 21x21 (:160-165)
  extractFeatures(curr)    -> mask around the previous frame's features (:76-79), FAST detect (:82),
                              FeatureSet::bucketingFeatures at the "todo" of :88 (src/bucket.cpp:24-68)
and the first max_tracks bucketed detections become the tracks of the next step (same rule as svo_gpu_step).

impl = "cv2": OpenCV itself + the reference's bucket.cpp compiled unmodified (the reference's own CPU code path)
impl = "orc": the plain-C restatement oracle/svo_oracle.c (bit-exact twin of the CUDA kernels)
"""
import numpy as np

from . import oracle as O


class CpuFrontend:
    def __init__(self, params, impl="cv2"):
        """params: dict with the fields of svo_gpu_step_params."""
        self.p = dict(params)
        self.impl = impl
        if impl == "cv2" and not O.have_cv2():
            raise RuntimeError("cv2 not importable")
        self.prev = None
        self.tracks = np.zeros((0, 2), np.float32)
        self.kf = True        # carry_tracks mode: the frame to come is a key-frame (a sequence's first frame always is)

    def reset(self):
        self.prev = None
        self.tracks = np.zeros((0, 2), np.float32)
        self.kf = True

    def step(self, img):
        p = self.p
        h, w = img.shape
        out = {}
        kw = dict(win=(p["win_w"], p["win_h"]), max_level=p["max_level"], max_count=p["max_count"],
                  epsilon=p["epsilon"], criteria_type=p["criteria_type"], flags=p["flags"],
                  min_eig_threshold=p["min_eig_threshold"])
        tracked = self.prev is not None and p["max_tracks"] > 0
        if tracked:
            klt = O.cv2_klt if self.impl == "cv2" else O.orc_klt
            out["next_xy"], out["status"], out["err"] = klt(self.prev, img, self.tracks, **kw)
        else:
            out["next_xy"], out["status"], out["err"] = (np.zeros((0, 2), np.float32), np.zeros(0, np.uint8),
                                                         np.zeros(0, np.float32))
        carry = bool(p.get("carry_tracks", 0))
        kf = self.kf if carry else True
        out["keyframe"] = int(kf)
        if kf:
            mask = None
            if p["mask_half"] > 0:
                pts = self.tracks if tracked else np.zeros((0, 2), np.float32)   # the points LK started from (src/tracking.cpp:77)
                mask = (O.cv2_build_mask if self.impl == "cv2" else O.orc_build_mask)(w, h, pts, p["mask_half"])
            fast = O.cv2_fast if self.impl == "cv2" else O.orc_fast
            kxy, kresp = fast(img, p["fast_threshold"], bool(p["fast_nms"]), mask)
            bucket = O.ref_bucket if (self.impl == "cv2" and O.have_ref_bucket()) else O.orc_bucket
            bxy, _ = bucket(kxy, np.zeros(len(kxy), np.int32), w, h, p["bucket_size"], p["per_bucket"])
        else:   # src/tracking.cpp:247: extractFeatures only runs on a key-frame
            kxy, kresp, bxy = np.zeros((0, 2), np.float32), np.zeros(0, np.float32), np.zeros((0, 2), np.float32)
        out["kp_xy"], out["kp_response"], out["bucket_xy"] = kxy, kresp, bxy
        if carry:
            # src/tracking.cpp:168-177: the survivors, in input order, are the frame's features; :254: a key-frame's new
            # features are appended; :69-70: the next frame is a key-frame iff this one is not and holds < features_to_track
            idx = np.flatnonzero(out["status"])
            out["survivor_index"] = idx.astype(np.int32)
            surv = out["next_xy"][idx]
            new = bxy[:max(0, p["max_tracks"] - len(surv))]
            self.tracks = np.ascontiguousarray(np.concatenate([surv, new]).astype(np.float32))
            self.kf = (not kf) and len(self.tracks) < p["features_to_track"]
        else:
            self.tracks = bxy[:p["max_tracks"]].copy()
        self.prev = img
        return out


def default_params(w, h, fast_threshold=20, bucket_size=32, per_bucket=5, max_tracks=2000, win=21, max_level=3,
                   mask_half=10.0, carry_tracks=0, features_to_track=70):
    """The reference's settings: FAST 20 + NMS (include/config_reader.h:35-38), LK 21x21, maxLevel 3,
    {COUNT+EPS, 50, 0.001}, OPTFLOW_LK_GET_MIN_EIGENVALS (src/tracking.cpp:157-165)."""
    return dict(w=w, h=h, fast_threshold=fast_threshold, fast_nms=1, mask_half=mask_half, bucket_size=bucket_size,
                per_bucket=per_bucket, max_tracks=max_tracks, win_w=win, win_h=win, max_level=max_level, max_count=50,
                epsilon=0.001, criteria_type=3, flags=8, min_eig_threshold=1e-4, carry_tracks=carry_tracks,
                features_to_track=features_to_track)
