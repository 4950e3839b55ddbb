"""Seeded synthetic frame source (no file I/O) standing in for the reference's camera feed
(/root/reference/src/lib/Camera.cpp:95-106 `GrabFrames`, out of scope) — SURVEY.md §8(d).

A "scene" is a textured canvas (multi-octave value noise + filled rectangles/discs that give true
corners).  Frames of a stream are crops of the canvas along a smooth camera path plus light sensor
noise, so consecutive frames share most of their features and the ratio test has survivors.
Pure numpy (float64 ufuncs only) => bit-identical across hosts for a given numpy version.
"""
import numpy as np


def _upsample_bilinear(g, H, W):
    gy, gx = g.shape
    ys = np.linspace(0.0, gy - 1.000001, H)
    xs = np.linspace(0.0, gx - 1.000001, W)
    y0 = ys.astype(np.int64); fy = (ys - y0)[:, None]
    x0 = xs.astype(np.int64); fx = (xs - x0)[None, :]
    rows = g[y0] * (1.0 - fy) + g[y0 + 1] * fy
    return rows[:, x0] * (1.0 - fx) + rows[:, x0 + 1] * fx


def make_canvas(H, W, seed=0, shapes_per_mpx=900, noise_cells=(96, 48, 24, 12, 6), contrast=1.0):
    """uint8 canvas HxW.  shapes_per_mpx tunes the corner density (report corner counts)."""
    rng = np.random.default_rng(seed)
    img = np.zeros((H, W), np.float64)
    amp = 1.0
    for c in noise_cells:
        g = rng.random((H // c + 3, W // c + 3))
        img += amp * _upsample_bilinear(g, H, W)
        amp *= 0.55
    img -= img.min()
    img *= 200.0 / max(img.max(), 1e-9)
    img += 25.0
    n_shapes = int(shapes_per_mpx * H * W / 1e6)
    xs = rng.integers(0, W, n_shapes); ys = rng.integers(0, H, n_shapes)
    ws = rng.integers(6, 48, n_shapes); hs = rng.integers(6, 48, n_shapes)
    delta = rng.integers(30, 110, n_shapes) * rng.choice([-1.0, 1.0], n_shapes) * contrast
    kind = rng.random(n_shapes)
    for i in range(n_shapes):
        x0, y0 = int(xs[i]), int(ys[i]); x1, y1 = min(W, x0 + int(ws[i])), min(H, y0 + int(hs[i]))
        if kind[i] < 0.7:
            img[y0:y1, x0:x1] += delta[i]
        else:
            r = 0.5 * min(x1 - x0, y1 - y0)
            if r < 2:
                continue
            yy, xx = np.mgrid[y0:y1, x0:x1]
            m = (yy - (y0 + y1 - 1) * 0.5) ** 2 + (xx - (x0 + x1 - 1) * 0.5) ** 2 <= r * r
            img[y0:y1, x0:x1][m] += delta[i]
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def camera_path(n, pad, seed=0):
    """Integer (ox, oy) crop offsets following a slow Lissajous path inside [0, pad]."""
    t = np.arange(n, dtype=np.float64)
    ph = (seed % 97) * 0.1
    ox = np.rint(pad * 0.5 * (1.0 + np.sin(0.031 * t + ph))).astype(np.int64)
    oy = np.rint(pad * 0.5 * (1.0 + np.sin(0.023 * t + 1.3 + ph))).astype(np.int64)
    return ox, oy


def make_stream(n, H=720, W=1280, seed=0, pad=96, sensor_noise=2, **canvas_kw):
    """uint8 [n, H, W] frames: crops of one canvas along camera_path + per-frame integer noise."""
    canvas = make_canvas(H + pad, W + pad, seed, **canvas_kw)
    ox, oy = camera_path(n, pad, seed)
    rng = np.random.default_rng(seed + 7919)
    out = np.empty((n, H, W), np.uint8)
    for i in range(n):
        f = canvas[oy[i]:oy[i] + H, ox[i]:ox[i] + W].astype(np.int16)
        if sensor_noise:
            f = f + rng.integers(-sensor_noise, sensor_noise + 1, f.shape, dtype=np.int16)
        out[i] = np.clip(f, 0, 255).astype(np.uint8)
    return out


def make_pair(H=480, W=640, seed=0, **kw):
    """Config 1 (BASELINE.json configs[0]): a frame pair with small camera motion."""
    s = make_stream(12, H, W, seed, **kw)
    return s[0], s[11]


def periodic_frame(H=480, W=640, period=37, seed=0):
    """Tie-stress image (SURVEY Appendix A `ties.py`): a tile repeated so scores/responses tie."""
    tile = make_canvas(period, period, seed, shapes_per_mpx=8000, noise_cells=(12, 6))
    reps = (H // period + 1, W // period + 1)
    return np.tile(tile, reps)[:H, :W].copy()


def random_descriptors(n, seed=0):
    return np.random.default_rng(seed).integers(0, 256, (n, 32), dtype=np.uint8)
