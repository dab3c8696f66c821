"""Seeded synthetic low-light images (SURVEY.md section 8(d) 'synthetic input').

Self-written test infrastructure; not code from the reference (which has none).
"""
from __future__ import annotations

import numpy as np


def synth_lowlight(h: int, w: int, seed: int = 0) -> np.ndarray:
    """float32 (h, w, 3) in [0, 1]; mean about 0.025, max about 0.13 (genuinely dark)."""
    rng = np.random.default_rng(seed)
    yy = np.linspace(0.0, 1.0, h, dtype=np.float64)[:, None]
    xx = np.linspace(0.0, 1.0, w, dtype=np.float64)[None, :]
    base = np.empty((h, w, 3), dtype=np.float64)
    for c in range(3):
        acc = np.zeros((h, w), dtype=np.float64)
        for _ in range(6):
            fy, fx = rng.uniform(0.5, 8.0, size=2)
            amp = rng.uniform(0.3, 1.0)
            py, px = rng.uniform(0.0, 2.0 * np.pi, size=2)
            acc += amp * np.sin(2 * np.pi * fy * yy + py) * np.sin(2 * np.pi * fx * xx + px)
        acc -= acc.min()
        acc /= max(acc.max(), 1e-12)
        base[:, :, c] = acc
    for _ in range(20):
        y0, x0 = int(rng.integers(0, h)), int(rng.integers(0, w))
        hh = int(rng.integers(max(2, h // 16), max(3, h // 3)))
        ww = int(rng.integers(max(2, w // 16), max(3, w // 3)))
        col = rng.uniform(0.0, 1.0, size=3)
        base[y0:y0 + hh, x0:x0 + ww, :] = 0.5 * base[y0:y0 + hh, x0:x0 + ww, :] + 0.5 * col
    x = 0.12 * base ** 2.2
    x = x + rng.normal(0.0, 0.01, size=x.shape) * np.sqrt(x) + rng.normal(0.0, 0.002, size=x.shape)
    return np.clip(x, 0.0, 1.0).astype(np.float32)
