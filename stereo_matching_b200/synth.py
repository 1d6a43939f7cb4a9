"""Synthetic stereo pairs with known ground truth (KITTI is not available offline).

`random_dot_pair` follows the generator SURVEY.md section 8(c) pins the known-answer hashes on:
``std::mt19937(seed)`` raw outputs ``& 255`` fill the RIGHT image in row-major order, the ground
truth disparity is `fg` inside the centre rectangle (W/3 < x < 2W/3, H/4 < y < 3H/4) and `bg`
elsewhere, and ``L(y, x) = R(y, max(x - d(y, x), 0))``.  numpy's legacy MT19937 seeding is the same
``init_genrand`` as ``std::mt19937(seed)``, so C++ and Python agree bit for bit.
"""
from __future__ import annotations

import numpy as np

FNV_OFFSET = 1469598103934665603
FNV_PRIME = 1099511628211


def _mt19937_bytes(seed: int, n: int) -> np.ndarray:
    bg = np.random.MT19937()
    bg._legacy_seeding(seed)  # == std::mt19937(seed)
    return (bg.random_raw(n) & 255).astype(np.uint8)


def ground_truth(h: int, w: int, bg: int = 12, fg: int = 40) -> np.ndarray:
    y, x = np.mgrid[0:h, 0:w]
    # the C++ form uses integer divisions: W/3 < x && x < 2*W/3 && H/4 < y && y < 3*H/4
    inside = (x > w // 3) & (x < 2 * w // 3) & (y > h // 4) & (y < 3 * h // 4)
    return np.where(inside, fg, bg).astype(np.int32)


def random_dot_pair(h: int, w: int, seed: int = 1234, bg: int = 12, fg: int = 40, noise: int = 0):
    """Returns (left, right, gt) as C-contiguous uint8/uint8/int32 arrays of shape (h, w).

    noise > 0 adds uniform integer noise in [-noise, noise] to the left image (clipped), so that
    uniqueness / LR invalidation and the hole-filling median actually fire.
    """
    right = _mt19937_bytes(seed, h * w).reshape(h, w)
    gt = ground_truth(h, w, bg, fg)
    xs = np.maximum(np.arange(w)[None, :] - gt, 0)
    left = np.take_along_axis(right, xs, axis=1)
    if noise > 0:
        rng = np.random.RandomState(seed ^ 0x5EED)
        n = rng.randint(-noise, noise + 1, size=(h, w))
        left = np.clip(left.astype(np.int32) + n, 0, 255).astype(np.uint8)
    return np.ascontiguousarray(left), np.ascontiguousarray(right), gt


def slanted_pair(h: int, w: int, seed: int = 7, d0: float = 8.0, d1: float = 40.0):
    """Slanted plane: disparity grows linearly with the row (fractional ground truth); the left image is
    a linear resampling of the right one.  Exercises the sub-pixel parabola fit."""
    right = _mt19937_bytes(seed, h * w).reshape(h, w)
    # smooth the dots a little so that linear resampling keeps texture
    r = right.astype(np.float64)
    r = (r + np.roll(r, 1, axis=1) + np.roll(r, -1, axis=1)) / 3.0
    right = np.clip(np.rint(r), 0, 255).astype(np.uint8)
    gt = (d0 + (d1 - d0) * np.arange(h)[:, None] / max(h - 1, 1)) * np.ones((1, w))
    xs = np.arange(w)[None, :] - gt
    x0 = np.floor(xs).astype(np.int64)
    fr = xs - x0
    x0c = np.clip(x0, 0, w - 1)
    x1c = np.clip(x0 + 1, 0, w - 1)
    rr = right.astype(np.float64)
    left = (1 - fr) * np.take_along_axis(rr, x0c, axis=1) + fr * np.take_along_axis(rr, x1c, axis=1)
    left = np.clip(np.rint(left), 0, 255).astype(np.uint8)
    return np.ascontiguousarray(left), np.ascontiguousarray(right), gt.astype(np.float32)


def sky_masks(h: int, w: int, frac: float = 0.25):
    """Synthetic sky masks (0/255) covering the top `frac` of both views (SURVEY 8(d) extra input iii)."""
    m = np.zeros((h, w), np.uint8)
    m[: int(h * frac), :] = 255
    return m, m.copy()


def fnv1a64(buf) -> str:
    """FNV-1a 64-bit over the raw bytes of `buf` (the hash SURVEY.md 8(c) quotes)."""
    data = np.ascontiguousarray(buf).view(np.uint8).ravel()
    hsh = FNV_OFFSET
    mask = (1 << 64) - 1
    for b in data.tobytes():
        hsh = ((hsh ^ b) * FNV_PRIME) & mask
    return f"{hsh:016x}"
