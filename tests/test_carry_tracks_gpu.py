"""svo_gpu_step with carry_tracks: the reference's own loop (src/tracking.cpp:240-268) instead of re-seeding every frame --
LK survivors (status != 0, input order, :168-175) carry over, detection (mask around the points LK started from :77, FAST :82,
bucketing :88) runs only on key-frames (a lane's first frame, or after a non-key-frame that ended below features_to_track,
:69-70), new features are appended after the survivors (:254).  The key-frame decision is taken per lane ON THE DEVICE.
Bit-exact per lane and per frame against the CPU pipeline built from the oracle (tracks, survivor indices, key-frame flags,
keypoints, bucketed features), with two steps outstanding."""
import numpy as np
import pytest

from svo_b200 import capi, synth
from svo_b200.frontend import Frontend

pytestmark = pytest.mark.gpu

W, H = 320, 240


def _run(oracle, features_to_track, T, max_tracks, depth, seeds=(5, 6, 7), speed=(1.0, 1.0, 1.0)):
    import torch
    from oracle.cpu_pipeline import CpuFrontend
    seqs = np.stack([synth.sequence(W, H, s, T, times=[v * t for t in range(T)]) for s, v in zip(seeds, speed)])   # [lane, t, h, w]
    B = len(seeds)
    frames = torch.from_numpy(np.ascontiguousarray(seqs.transpose(1, 0, 2, 3))).pin_memory()
    kfs = []
    with capi.Context(max_w=W, max_h=H, max_levels=4, max_points=8192, max_batch=B) as ctx:
        fe = Frontend(ctx, W, H, B, kp_capacity=8192, bucket_capacity=2048, max_tracks=max_tracks, mask_half=10.0,
                      carry_tracks=1, features_to_track=features_to_track)
        cpus = [CpuFrontend(fe.pdict, impl="orc") for _ in range(B)]
        done = 0

        def check():
            nonlocal done
            fe.wait()
            t = done
            done += 1
            row = []
            for b in range(B):
                ref = cpus[b].step(seqs[b, t])
                nt = int(fe.n_tracked[b])
                assert nt == len(ref["next_xy"]), (t, b)
                assert np.array_equal(fe.status[b, :nt], ref["status"]) and np.array_equal(fe.next_xy[b, :nt], ref["next_xy"])
                assert np.array_equal(fe.err[b, :nt], ref["err"])
                assert int(fe.keyframe[b]) == ref["keyframe"], (t, b)
                ns = int(fe.n_survived[b])
                assert ns == len(ref["survivor_index"]) and np.array_equal(fe.survivor_index[b, :ns], ref["survivor_index"])
                nk, nb = int(fe.n_keypoints[b]), int(fe.n_bucketed[b])
                assert nk == len(ref["kp_xy"]) and nb == len(ref["bucket_xy"]), (t, b, nk, nb)
                assert np.array_equal(fe.kp_xy[b, :nk], ref["kp_xy"]) and np.array_equal(fe.kp_response[b, :nk], ref["kp_response"])
                assert np.array_equal(fe.bucket_xy[b, :nb], ref["bucket_xy"])
                row.append(ref["keyframe"])
            kfs.append(row)

        for t in range(T):
            fe.submit(frames[t].data_ptr(), frames[t + 1].data_ptr() if t + 1 < T else None, capi.MEM_HOST, W, W * H)
            if fe.outstanding == depth:
                check()
        while fe.outstanding:
            check()
    return np.array(kfs)


@pytest.mark.parametrize("depth", [1, 2])
def test_reference_loop_with_the_shipped_threshold(oracle, depth):
    """features_to_track = 70 (configs/config.yaml:15): after the first frame the tracks simply carry over."""
    kf = _run(oracle, 70, 6, 400, depth)
    assert kf[0].all() and not kf[1:].any()


def test_key_frames_fire_per_lane(oracle):
    """A threshold at the track count and lanes that move at different speeds (6, 0.2 and 3 px per frame): the fast lanes
    lose tracks at the image border and re-detect, the slow one does not, so key-frames fire in some lanes and not in others
    within the same launch; a key-frame is never followed directly by another (src/tracking.cpp:69)."""
    kf = _run(oracle, 400, 10, 400, 2, speed=(3.0, 0.1, 1.5))
    assert kf[0].all() and kf[1:].any() and not kf[1:].all()
    assert (kf[1:].min(axis=1) != kf[1:].max(axis=1)).any()          # lanes disagree within one step
    assert not (kf[1:] & kf[:-1]).any()
