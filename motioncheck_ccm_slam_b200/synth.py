"""Seeded synthetic grayscale frame streams for parity tests and benchmarks.

A stream is one static scene (multi-octave value noise + filled rectangles/discs for strong FAST
corners) viewed through a window that moves 2 px right / 1 px down per frame, plus fresh ±4 pixel
noise per frame, so consecutive frames share most of their corners. numpy only.
"""
import numpy as np


def _value_noise(rng, h, w, cell, amp):
    gh, gw = h // cell + 2, w // cell + 2
    g = rng.uniform(-amp, amp, size=(gh, gw)).astype(np.float32)
    ys = np.arange(h, dtype=np.float32) / cell
    xs = np.arange(w, dtype=np.float32) / cell
    y0 = ys.astype(np.int32)
    x0 = xs.astype(np.int32)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    a = g[y0][:, x0]
    b = g[y0][:, x0 + 1]
    c = g[y0 + 1][:, x0]
    d = g[y0 + 1][:, x0 + 1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def make_scene(seed, h, w, n_shapes=None):
    """Static scene canvas, float32 in [0,255]."""
    rng = np.random.default_rng(seed)
    img = np.full((h, w), 128.0, dtype=np.float32)
    for cell, amp in ((64, 60.0), (16, 40.0), (4, 20.0)):
        img += _value_noise(rng, h, w, cell, amp)
    if n_shapes is None:
        n_shapes = max(200, (h * w) // 1800)
    yy, xx = np.mgrid[0:h, 0:w]
    for _ in range(n_shapes):
        cx, cy = int(rng.integers(0, w)), int(rng.integers(0, h))
        val = float(rng.integers(0, 256))
        if rng.random() < 0.5:
            rw, rh = int(rng.integers(4, 40)), int(rng.integers(4, 40))
            img[max(0, cy - rh):cy + rh, max(0, cx - rw):cx + rw] = val
        else:
            r = int(rng.integers(3, 24))
            y0, y1, x0, x1 = max(0, cy - r), min(h, cy + r + 1), max(0, cx - r), min(w, cx + r + 1)
            m = (yy[y0:y1, x0:x1] - cy) ** 2 + (xx[y0:y1, x0:x1] - cx) ** 2 <= r * r
            img[y0:y1, x0:x1][m] = val
    return np.clip(img, 0, 255)


def make_stream(seed, n_frames, w=752, h=480):
    """uint8 [n_frames, h, w]; frame t is the scene window at offset (2t, t) plus fresh ±4 noise."""
    scene = make_scene(seed, h + n_frames, w + 2 * n_frames)
    out = np.empty((n_frames, h, w), dtype=np.uint8)
    for t in range(n_frames):
        rng = np.random.default_rng(seed * 7919 + t + 1)
        f = scene[t:t + h, 2 * t:2 * t + w] + rng.integers(-4, 5, size=(h, w)).astype(np.float32)
        out[t] = np.clip(np.rint(f), 0, 255).astype(np.uint8)
    return out


def make_frame(seed, w=752, h=480):
    return make_stream(seed, 1, w, h)[0]
