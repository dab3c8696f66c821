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


# Content classes of hard_cases(): "bw" = pure black (x = 0, I = 1) beside pure white (x = 1, I = 256), the largest
# contrast the operator accepts; "full" = full-range content without that adjacency; "dark" = low-light range.
HARD_CASE_CLASS = {
    "edge_v_0_1": "bw", "edge_h_0_1": "bw", "bar16_white_on_black": "bw", "line1_white_on_black": "bw",
    "corner_coloured": "bw", "edge_v_lowlight": "dark", "edge_v_1_255": "full", "stripes64": "full",
    "stripes256_h": "full", "noise_uniform": "full", "synth_x7p8_clipped": "full",
}


def hard_cases(h: int, w: int) -> dict:
    """Hard content for the surround blur: name -> float32 (h, w, 3) in [0, 1].

    Straight full-range edges (the oracle's hard 4-sigma cut-off shows 4 sigma into the dark side), a 16-px bar and
    a 1-px line of white on black, 64- / 256-px 0/1 stripes, uniform noise on [0, 1], the synthetic frame x 7.8
    (clipped), a low-light edge and an edge between 1/255 and 1."""
    rng = np.random.default_rng(7)

    def z(v=0.0):
        return np.full((h, w, 3), v, np.float32)

    c = {}
    a = z(); a[:, w // 2:] = 1.0; c["edge_v_0_1"] = a
    a = z(); a[h // 2:, :] = 1.0; c["edge_h_0_1"] = a
    a = z(0.004); a[:, w // 2:] = 0.13; c["edge_v_lowlight"] = a
    a = z(1.0 / 255.0); a[:, w // 2:] = 1.0; c["edge_v_1_255"] = a
    a = z(); a[:, w // 2 - 8:w // 2 + 8] = 1.0; c["bar16_white_on_black"] = a
    a = z(); a[:, w // 2] = 1.0; c["line1_white_on_black"] = a
    a = z(); a[:, (np.arange(w) // 64) % 2 == 1] = 1.0; c["stripes64"] = a
    a = z(); a[(np.arange(h) // 256) % 2 == 1, :] = 1.0; c["stripes256_h"] = a
    c["noise_uniform"] = rng.random((h, w, 3), dtype=np.float32)
    a = z(); a[h // 3:, w // 3:] = 1.0; a[:, :, 1] *= 0.5; c["corner_coloured"] = a
    c["synth_x7p8_clipped"] = np.clip(synth_lowlight(h, w, 3) * np.float32(7.8), 0.0, 1.0).astype(np.float32)
    assert set(c) == set(HARD_CASE_CLASS)
    return c
