"""Deterministic synthetic video source (numpy only; no cv2, no GPU).

Used by bench.py, the tests and the golden-vector script to produce the frames
BASELINE.json's configs are quoted on ("synthetic 640x480", "1920x1080 synthetic
video", ...).  It follows the plan of SURVEY.md section 8(d): a textured canvas
(low-frequency base + random filled shapes + slight blur) that is panned by
(2, 1) px per frame with fresh sensor noise, so consecutive frames match well and
every pyramid level fills its ORB quota.
"""
import numpy as np

MARGIN = 64


def _upsample_bilinear(g, h, w):
    gh, gw = g.shape
    ys = np.linspace(0, gh - 1, h)
    xs = np.linspace(0, gw - 1, w)
    y0 = np.floor(ys).astype(int).clip(0, gh - 2)
    x0 = np.floor(xs).astype(int).clip(0, gw - 2)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    a = g[y0][:, x0]
    b = g[y0][:, x0 + 1]
    c = g[y0 + 1][:, x0]
    d = g[y0 + 1][:, x0 + 1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def _blur(img, sigma=0.8):
    r = 2
    k = np.exp(-0.5 * (np.arange(-r, r + 1) / sigma) ** 2)
    k /= k.sum()
    p = np.pad(img, ((0, 0), (r, r)), mode="reflect")
    img = sum(k[i] * p[:, i:i + img.shape[1]] for i in range(2 * r + 1))
    p = np.pad(img, ((r, r), (0, 0)), mode="reflect")
    return sum(k[i] * p[i:i + img.shape[0], :] for i in range(2 * r + 1))


def canvas(w, h, seed=0):
    """float32 canvas of size (h + 2*MARGIN, w + 2*MARGIN)."""
    rng = np.random.default_rng(seed)
    ch, cw = h + 2 * MARGIN, w + 2 * MARGIN
    base = rng.integers(0, 256, (ch // 32 + 2, cw // 32 + 2)).astype(np.float64)
    img = _upsample_bilinear(base, ch, cw) * 0.5 + 64.0
    n_shapes = int(cw * ch / 600)
    kinds = rng.integers(0, 3, n_shapes)
    cxs = rng.integers(0, cw, n_shapes)
    cys = rng.integers(0, ch, n_shapes)
    sizes = rng.integers(4, 41, n_shapes)
    grays = rng.integers(0, 256, n_shapes)
    for kind, cx, cy, s, g in zip(kinds, cxs, cys, sizes, grays):
        x0, x1 = max(cx - s, 0), min(cx + s + 1, cw)
        y0, y1 = max(cy - s, 0), min(cy + s + 1, ch)
        if x0 >= x1 or y0 >= y1:
            continue
        yy, xx = np.mgrid[y0:y1, x0:x1]
        if kind == 0:      # axis-aligned rectangle, half extents (s, s*0.6)
            m = (np.abs(xx - cx) <= s) & (np.abs(yy - cy) <= int(s * 0.6))
        elif kind == 1:    # disc
            m = (xx - cx) ** 2 + (yy - cy) ** 2 <= (s // 2 + 2) ** 2
        else:              # right triangle
            m = (xx - cx >= 0) & (yy - cy >= 0) & ((xx - cx) + (yy - cy) <= s)
        img[y0:y1, x0:x1][m] = g
    return _blur(img, 0.8).astype(np.float32)


class SyntheticVideo:
    """frame(k) -> uint8[h, w]; deterministic in (w, h, seed, k)."""

    def __init__(self, w=640, h=480, seed=0):
        self.w, self.h, self.seed = w, h, seed
        self._canvas = canvas(w, h, seed)

    def frame(self, k):
        ox = MARGIN + (2 * k) % MARGIN
        oy = MARGIN + k % MARGIN
        crop = self._canvas[oy:oy + self.h, ox:ox + self.w]
        rng = np.random.default_rng(self.seed * 1000003 + k + 1)
        noisy = crop + rng.normal(0.0, 2.0, crop.shape).astype(np.float32)
        return np.clip(np.rint(noisy), 0, 255).astype(np.uint8)

    def frames(self, start, count):
        return np.stack([self.frame(start + i) for i in range(count)])
