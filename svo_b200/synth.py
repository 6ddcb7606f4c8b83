"""Seeded synthetic image sequences (numpy only; no dataset, no cv2) -- SURVEY.md section 8(d).

A canvas of (H+128)x(W+128) pixels = 0.6 * cubic-upsampled(x4) uniform noise + 0.4 * Gaussian-blurred (sigma 1.2)
uniform noise; frame t is a crop of that canvas under a similarity warp (scale 1+0.004t, rotation 0.001t rad,
translation (-1.7t, +0.9t) px), sampled bilinearly.  About 2 px/frame of motion, textured enough that FAST(20) yields
~10k raw corners at 1241x376 and essentially every track is trackable.

The workload table WORKLOADS mirrors BASELINE.json's configs C1..C4 (C5 = 64 x C1).
"""
import numpy as np

MARGIN = 64

# name -> dict(w, h, n_tracks, fast_thr, bucket_size, per_bucket, win, max_level, seed)
WORKLOADS = {
    "C1": dict(w=1241, h=376, n_tracks=2000, fast_thr=20, bucket_size=32, per_bucket=5, win=21, max_level=3, seed=11),
    "C2": dict(w=752, h=480, n_tracks=5000, fast_thr=10, bucket_size=24, per_bucket=8, win=21, max_level=3, seed=12),
    "C3": dict(w=1920, h=1080, n_tracks=10000, fast_thr=20, bucket_size=32, per_bucket=5, win=21, max_level=4, seed=13),
    "C4": dict(w=3840, h=2160, n_tracks=50000, fast_thr=20, bucket_size=24, per_bucket=8, win=31, max_level=5, seed=14),
    # C1 with the reference's stereo-matching window (src/tracking.cpp:104: 11x11): times the 11x11 LK instantiation
    "C1w11": dict(w=1241, h=376, n_tracks=2000, fast_thr=20, bucket_size=32, per_bucket=5, win=11, max_level=3, seed=11),
}


def _cubic_weights(frac, a=-0.75):
    """Keys cubic convolution weights for taps at -1, 0, 1, 2 given the fractional position."""
    f = frac[:, None] + np.array([1.0, 0.0, -1.0, -2.0])[None, :]
    f = np.abs(f)
    w = np.where(f <= 1.0, (a + 2) * f ** 3 - (a + 3) * f ** 2 + 1,
                 np.where(f < 2.0, a * f ** 3 - 5 * a * f ** 2 + 8 * a * f - 4 * a, 0.0))
    return w


def _upsample_axis(a, factor, axis):
    n = a.shape[axis]
    o = np.arange(n * factor, dtype=np.float64)
    s = (o + 0.5) / factor - 0.5
    i0 = np.floor(s).astype(np.int64)
    w = _cubic_weights(s - i0)
    out = 0.0
    for k in range(4):
        idx = np.clip(i0 + k - 1, 0, n - 1)
        shape = [1] * a.ndim
        shape[axis] = -1
        out = out + np.take(a, idx, axis=axis) * w[:, k].reshape(shape)
    return out


def _blur_axis(a, sigma, axis):
    r = int(np.ceil(3 * sigma))
    x = np.arange(-r, r + 1, dtype=np.float64)
    k = np.exp(-0.5 * (x / sigma) ** 2)
    k /= k.sum()
    pad = [(0, 0)] * a.ndim
    pad[axis] = (r, r)
    p = np.pad(a, pad, mode="reflect")
    n = a.shape[axis]
    out = 0.0
    for i in range(2 * r + 1):
        sl = [slice(None)] * a.ndim
        sl[axis] = slice(i, i + n)
        out = out + p[tuple(sl)] * k[i]
    return out


def make_canvas(w, h, seed):
    """float32 canvas of shape (h + 2*MARGIN, w + 2*MARGIN), values in [0, 255]."""
    rng = np.random.default_rng(seed)
    ch, cw = h + 2 * MARGIN, w + 2 * MARGIN
    lh, lw = (ch + 3) // 4, (cw + 3) // 4
    low = rng.integers(0, 256, size=(lh, lw)).astype(np.float64)
    up = _upsample_axis(_upsample_axis(low, 4, 0), 4, 1)[:ch, :cw]
    fine = rng.integers(0, 256, size=(ch, cw)).astype(np.float64)
    fine = _blur_axis(_blur_axis(fine, 1.2, 0), 1.2, 1)
    return np.clip(0.6 * up + 0.4 * fine, 0, 255).astype(np.float32)


def render_frame(canvas, w, h, t):
    """uint8 frame (h, w) at (possibly fractional) time t."""
    t = float(t)
    s = 1.0 + 0.004 * t
    th = 0.001 * t
    tx, ty = -1.7 * t, 0.9 * t
    ys, xs = np.mgrid[0:h, 0:w].astype(np.float32)
    cx, cy = np.float32(w / 2.0), np.float32(h / 2.0)
    dx, dy = xs - cx, ys - cy
    c, sn = np.float32(np.cos(th) * s), np.float32(np.sin(th) * s)
    u = c * dx - sn * dy + cx + np.float32(MARGIN + tx)
    v = sn * dx + c * dy + cy + np.float32(MARGIN + ty)
    ch, cw = canvas.shape
    u = np.clip(u, 0, cw - 1.001)
    v = np.clip(v, 0, ch - 1.001)
    u0 = np.floor(u).astype(np.int64)
    v0 = np.floor(v).astype(np.int64)
    fu = u - u0
    fv = v - v0
    flat = canvas.reshape(-1)
    i00 = v0 * cw + u0
    val = (flat[i00] * (1 - fu) * (1 - fv) + flat[i00 + 1] * fu * (1 - fv) +
           flat[i00 + cw] * (1 - fu) * fv + flat[i00 + cw + 1] * fu * fv)
    return np.clip(np.rint(val), 0, 255).astype(np.uint8)


def sequence(w, h, seed, n_frames, times=None):
    """uint8 array [n_frames, h, w]; times defaults to 0..n_frames-1."""
    canvas = make_canvas(w, h, seed)
    if times is None:
        times = range(n_frames)
    return np.stack([render_frame(canvas, w, h, t) for t in times])


def ring_times(n):
    """A closed loop of n time stamps with unit steps everywhere, e.g. n=8 -> 0,1,2,3,4,3,2,1 (so that frame n-1 -> frame 0
    is also a ~2 px motion).  Used by bench.py to cycle through more frame data than fits in L2."""
    half = n // 2
    return [i if i <= half else n - i for i in range(n)]


def noise_image(w, h, seed):
    return np.random.default_rng(seed).integers(0, 256, size=(h, w), dtype=np.uint8)


def block_image(w, h, seed, block=8):
    """Piecewise-constant blocks: FAST + strict NMS returns no keypoint at all (ties suppress both), NMS-off returns many."""
    rng = np.random.default_rng(seed)
    small = rng.integers(0, 256, size=((h + block - 1) // block, (w + block - 1) // block), dtype=np.uint8)
    return np.kron(small, np.ones((block, block), np.uint8))[:h, :w].copy()
