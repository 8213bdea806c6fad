"""Synthetic luma sequences: the data generators shared by tests and bench."""
import numpy as np


def synth_sequence(width, height, n, seed=1, motion=(2, 1), noise=2.0, rects=16):
    """Config-2 style luma: band-limited noise translating `motion` px/frame,
    moving 64x64 rectangles, per-frame noise. Returns uint8 array [n, height, width]."""
    rng = np.random.default_rng(seed)
    big_h, big_w = height + 64 + abs(motion[1]) * n + 8, width + 64 + abs(motion[0]) * n + 8
    base = rng.integers(0, 256, size=(big_h // 8 + 2, big_w // 8 + 2)).astype(np.float32)
    base = np.kron(base, np.ones((8, 8), dtype=np.float32))[:big_h, :big_w]
    k = np.ones(5, dtype=np.float32) / 5
    base = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 1, base)
    base = np.apply_along_axis(lambda c: np.convolve(c, k, mode="same"), 0, base)
    fine = rng.integers(-24, 25, size=(big_h, big_w)).astype(np.float32)
    base = np.clip(base + fine, 0, 255)
    rx = rng.integers(0, max(1, width - 64), size=rects)
    ry = rng.integers(0, max(1, height - 64), size=rects)
    rv = rng.integers(-6, 7, size=(rects, 2))
    rc = rng.integers(0, 256, size=rects)
    out = np.empty((n, height, width), dtype=np.uint8)
    for t in range(n):
        ox, oy = 32 + max(0, motion[0]) * t if motion[0] >= 0 else 32 - motion[0] * (n - t), 0
        ox = 32 + (motion[0] * t if motion[0] >= 0 else -motion[0] * (n - 1 - t))
        oy = 32 + (motion[1] * t if motion[1] >= 0 else -motion[1] * (n - 1 - t))
        f = base[oy:oy + height, ox:ox + width].copy()
        for i in range(rects):
            x = int(np.clip(rx[i] + rv[i, 0] * t, 0, max(0, width - 64)))
            y = int(np.clip(ry[i] + rv[i, 1] * t, 0, max(0, height - 64)))
            f[y:y + 64, x:x + 64] = rc[i]
        if noise > 0:
            f = f + rng.normal(0, noise, size=f.shape).astype(np.float32)
        out[t] = np.clip(np.rint(f), 0, 255).astype(np.uint8)
    return out


def random_frames(width, height, n, seed=13596):
    """Config-1 style: i.i.d. uniform u8 frames (the SadTest RANDOM pattern)."""
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, size=(n, height, width), dtype=np.uint8)


