"""Seeded synthetic inputs of the shapes BASELINE.json names: corner-rich grayscale frames,
stereo partners with a piecewise-constant disparity field, and the map/frame structures the matchers consume.
Pure numpy; identical bytes for identical seeds on every machine (PCG64)."""
import numpy as np


def _value_noise(rng, h, w, cell, amp):
    gh, gw = h // cell + 2, w // cell + 2
    g = rng.random((gh, gw), dtype=np.float32)
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
    return amp * ((a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy)


def image(seed, w, h, n_rect=90, n_tri=50, noise_sigma=2.0):
    """I(seed, w, h): 4 octaves of bilinear value noise + filled rectangles and triangles + Gaussian noise."""
    rng = np.random.Generator(np.random.PCG64(seed))
    img = np.full((h, w), 38.0, np.float32)
    for cell, amp in ((64, 96.0), (32, 48.0), (16, 24.0), (8, 12.0)):
        img += _value_noise(rng, h, w, cell, amp)
    for _ in range(n_rect):
        x0, y0 = int(rng.integers(0, w - 8)), int(rng.integers(0, h - 8))
        rw, rh = int(rng.integers(6, w // 6)), int(rng.integers(6, h // 6))
        img[y0:y0 + rh, x0:x0 + rw] = float(rng.integers(0, 256))
    yy, xx = np.mgrid[0:h, 0:w]
    for _ in range(n_tri):
        cx, cy = rng.integers(0, w), rng.integers(0, h)
        p = np.stack([cx + rng.integers(-w // 10, w // 10, 3), cy + rng.integers(-h // 8, h // 8, 3)], 1).astype(np.float32)
        x0, x1 = int(max(p[:, 0].min(), 0)), int(min(p[:, 0].max() + 1, w))
        y0, y1 = int(max(p[:, 1].min(), 0)), int(min(p[:, 1].max() + 1, h))
        if x1 <= x0 or y1 <= y0:
            continue
        X = xx[y0:y1, x0:x1].astype(np.float32)
        Y = yy[y0:y1, x0:x1].astype(np.float32)

        def edge(a, b):
            return (b[0] - a[0]) * (Y - a[1]) - (b[1] - a[1]) * (X - a[0])

        e0, e1, e2 = edge(p[0], p[1]), edge(p[1], p[2]), edge(p[2], p[0])
        inside = ((e0 >= 0) & (e1 >= 0) & (e2 >= 0)) | ((e0 <= 0) & (e1 <= 0) & (e2 <= 0))
        img[y0:y1, x0:x1][inside] = float(rng.integers(0, 256))
    img += rng.normal(0.0, noise_sigma, (h, w)).astype(np.float32)
    return np.clip(img + 0.5, 0, 255).astype(np.uint8)


def stereo_right(left, seed, d_min=4.0, d_max=40.0, noise_sigma=1.0):
    """Right view: the left image shifted by a piecewise-constant disparity (3 depth planes laid out as vertical
    bands with seeded borders), bilinear in x, plus independent sensor noise."""
    rng = np.random.Generator(np.random.PCG64(seed))
    h, w = left.shape
    cuts = np.sort(rng.integers(w // 6, w - w // 6, 2))
    disp = np.empty(w, np.float32)
    d = rng.uniform(d_min, d_max, 3).astype(np.float32)
    disp[:cuts[0]] = d[0]
    disp[cuts[0]:cuts[1]] = d[1]
    disp[cuts[1]:] = d[2]
    xs = np.arange(w, dtype=np.float32) + disp  # right(x) = left(x + d)
    x0 = np.clip(np.floor(xs).astype(np.int32), 0, w - 1)
    x1 = np.clip(x0 + 1, 0, w - 1)
    fx = (xs - np.floor(xs))[None, :]
    L = left.astype(np.float32)
    r = L[:, x0] * (1 - fx) + L[:, x1] * fx
    r += rng.normal(0.0, noise_sigma, (h, w)).astype(np.float32)
    return np.ascontiguousarray(np.clip(r + 0.5, 0, 255).astype(np.uint8))


def adversarial(kind, w, h, seed=0):
    """Parity-only inputs: constant (no keypoints), 8-px checkerboard (response ties), uniform noise."""
    if kind == "constant":
        return np.full((h, w), 117, np.uint8)
    if kind == "checker":
        yy, xx = np.mgrid[0:h, 0:w]
        return (((yy // 8 + xx // 8) % 2) * 255).astype(np.uint8)
    if kind == "noise":
        return np.random.Generator(np.random.PCG64(seed)).integers(0, 256, (h, w), dtype=np.uint8)
    raise ValueError(kind)


SHAPES = {"euroc": (752, 480), "kitti": (1241, 376), "tum": (640, 480)}
