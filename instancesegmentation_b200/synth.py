"""Deterministic synthetic inputs shared by tests/, bench.py and the golden generator.

numpy ``default_rng`` only (PCG64 is platform independent), so the same arrays are
produced in the dev container and on the GPU box.  Images imitate what the reference
feeds the network (train_instance.py:80-85: ToTensor + Normalize(0.5, 0.5) -> [-1, 1]).
"""
import numpy as np


def _upsample_linear(a, H, W):
    """Separable linear resize of [h,w,c] to [H,W,c] (half-pixel centres)."""
    h, w, _ = a.shape
    ys = np.clip((np.arange(H) + 0.5) * h / H - 0.5, 0, h - 1)
    xs = np.clip((np.arange(W) + 0.5) * w / W - 0.5, 0, w - 1)
    y0 = np.floor(ys).astype(int); y1 = np.minimum(y0 + 1, h - 1); fy = (ys - y0)[:, None, None]
    x0 = np.floor(xs).astype(int); x1 = np.minimum(x0 + 1, w - 1); fx = (xs - x0)[None, :, None]
    rows = a[y0] * (1 - fy) + a[y1] * fy
    return rows[:, x0] * (1 - fx) + rows[:, x1] * fx


def image_u8(H, W, seed, channels=3):
    """'natural-like' uint8 image [H,W,C]: smooth noise + 6 filled ellipses."""
    rng = np.random.default_rng(seed)
    low = rng.uniform(0, 255, (max(H // 16, 2), max(W // 16, 2), channels))
    img = _upsample_linear(low, H, W)
    yy, xx = np.mgrid[0:H, 0:W]
    for _ in range(6):
        cy, cx = rng.uniform(0, H), rng.uniform(0, W)
        ry, rx = rng.uniform(H / 16, H / 3), rng.uniform(W / 16, W / 3)
        col = rng.uniform(0, 255, channels)
        inside = ((yy - cy) / ry) ** 2 + ((xx - cx) / rx) ** 2 <= 1.0
        img[inside] = 0.35 * img[inside] + 0.65 * col
    img += rng.normal(0, 2.0, img.shape)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def batch(B, H, W, seed=1234, channels=3, kind="natural"):
    """float32 NCHW batch in [-1,1]; image i uses seed ``seed + i``."""
    out = np.empty((B, channels, H, W), np.float32)
    for i in range(B):
        if kind == "noise":
            out[i] = np.random.default_rng(seed + i).uniform(-1, 1, (channels, H, W))
        else:
            u8 = image_u8(H, W, seed + i, channels)
            out[i] = (u8.transpose(2, 0, 1).astype(np.float32) / 255.0 - 0.5) / 0.5
    return out


def boxes(n_img, per_img, canvas_h, canvas_w, seed=7, lo=96, hi=640):
    """Instance boxes (x0,y0,wk,hk) int32 [n_img*per_img,4] inside the canvas, plus the CSR
    ``img_first`` [n_img+1] (instances of one image are contiguous)."""
    rng = np.random.default_rng(seed)
    hi_w, hi_h = min(hi, canvas_w), min(hi, canvas_h)
    lo_w, lo_h = min(lo, hi_w), min(lo, hi_h)
    out = np.empty((n_img * per_img, 4), np.int32)
    for i in range(n_img * per_img):
        wk = int(rng.integers(lo_w, hi_w + 1)); hk = int(rng.integers(lo_h, hi_h + 1))
        out[i] = (int(rng.integers(0, canvas_w - wk + 1)), int(rng.integers(0, canvas_h - hk + 1)), wk, hk)
    return out, (np.arange(n_img + 1) * per_img).astype(np.int32)


def logits(N, h, w, seed=11, scale=2.0):
    """Smooth, blob-like fp32 logit maps [N,h,w] with values crossing zero."""
    rng = np.random.default_rng(seed)
    out = np.empty((N, h, w), np.float32)
    for n in range(N):
        low = rng.normal(0, scale, (max(h // 32, 2), max(w // 32, 2), 1))
        out[n] = _upsample_linear(low, h, w)[..., 0] + rng.normal(0, 0.05, (h, w))
    return out
