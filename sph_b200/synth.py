"""Synthetic image stacks of the BASELINE.json shapes.

float32, point-major [H*W, d]; pixel id = row*W + col (sph/utils/ImageHelper.cpp:8-29).
Seeds are fixed so the CPU oracle and the GPU path see identical bytes.
"""
from __future__ import annotations

import numpy as np


def _blur(field: np.ndarray, sigma: float) -> np.ndarray:
    from scipy.ndimage import gaussian_filter

    return gaussian_filter(field, sigma=sigma, mode="reflect")


def hyperspectral(h: int, w: int, d: int, endmembers: int = 16, seed: int = 1001, blur_sigma: float = 6.0, noise: float = 0.01) -> np.ndarray:
    """M random endmember spectra (uniform, 5-tap box smoothed along d), abundances =
    softmax of M blurred N(0,1) fields, + N(0, noise²), clipped to [0,1]."""
    rng = np.random.default_rng(seed)
    spectra = rng.uniform(0.0, 1.0, size=(endmembers, d + 4))
    kernel = np.ones(5) / 5.0
    spectra = np.stack([np.convolve(s, kernel, mode="valid") for s in spectra]).astype(np.float32)  # [M, d]
    fields = np.empty((endmembers, h, w), np.float32)
    for m in range(endmembers):
        f = _blur(rng.standard_normal((h, w)).astype(np.float32), blur_sigma)
        fields[m] = f / (f.std() + 1e-12)
    fields *= 2.0
    fields -= fields.max(axis=0, keepdims=True)
    np.exp(fields, out=fields)
    fields /= fields.sum(axis=0, keepdims=True)
    ab = fields.reshape(endmembers, h * w).T  # [N, M]
    out = np.empty((h * w, d), np.float32)
    step = 1 << 18
    for s in range(0, h * w, step):
        e = min(h * w, s + step)
        blk = ab[s:e] @ spectra
        blk += rng.standard_normal(blk.shape).astype(np.float32) * noise
        np.clip(blk, 0.0, 1.0, out=blk)
        out[s:e] = blk
    return out


def rgb(h: int, w: int, seed: int = 1003, blur_sigma: float = 12.0, noise: float = 0.02) -> np.ndarray:
    """3 blurred uniform fields + noise, quantised to 1/255 steps (exact duplicates and ties on purpose)."""
    rng = np.random.default_rng(seed)
    out = np.empty((h * w, 3), np.float32)
    for c in range(3):
        f = _blur(rng.uniform(0.0, 1.0, (h, w)).astype(np.float32), blur_sigma)
        f = (f - f.min()) / (f.max() - f.min() + 1e-12)
        f += rng.standard_normal((h, w)).astype(np.float32) * noise
        q = np.clip(np.rint(f * 255.0), 0, 255).astype(np.float32) / np.float32(255.0)
        out[:, c] = q.reshape(-1)
    return out


def s_curve(n: int = 1000, noise: float = 0.2, seed: int = 123) -> np.ndarray:
    """S-curve manifold in 3-D (shape of sph/utils/TestData createSCurve; own generator, numpy RNG)."""
    rng = np.random.default_rng(seed)
    t = 3.0 * np.pi * (rng.uniform(size=n) - 0.5)
    x = np.sin(t)
    y = 2.0 * rng.uniform(size=n)
    z = np.sign(t) * (np.cos(t) - 1.0)
    pts = np.stack([x, y, z], axis=1) + noise * rng.standard_normal((n, 3))
    return pts.astype(np.float32)


CONFIGS = {
    # name: (h, w, d, generator kwargs)
    "C1": dict(h=145, w=145, d=200, endmembers=16, seed=1001),
    "C2": dict(h=512, w=217, d=204, endmembers=16, seed=1002),
    "C3": dict(h=1080, w=1920, d=3, seed=1003),
    "C4": dict(h=2048, w=2048, d=128, endmembers=32, seed=1004),
    "C5": dict(h=1024, w=1024, d=1000, endmembers=64, seed=1005),
}


def make(name: str) -> np.ndarray:
    cfg = dict(CONFIGS[name])
    if cfg["d"] == 3:
        return rgb(cfg["h"], cfg["w"], seed=cfg["seed"])
    return hyperspectral(**cfg)
