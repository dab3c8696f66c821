"""NumPy oracle for MSRCR Retinex (BASELINE.json north_star (a)).

Self-written oracle -- not code from CILAB-IITM/Low_Light_Image_Enhancement
(the reference holds no code: /root/reference/README.md:1-2).  Test
infrastructure only; float64 arithmetic on float32 inputs.

    I      = 255 x + 1
    k_s[t] = exp(-t^2 / (2 s^2)), t = -r..r, r = ceil(truncate * s), sum-normalised
    S_s    = column_blur(row_blur(I, k_s), k_s), index clamp at the borders
    R      = sum_s w_s (ln I - ln S_s)
    C      = beta (ln(alpha I) - ln(I_R + I_G + I_B))
    out    = clip((G C R + b) / 255, 0, 1)
    finish: instead of the plain clip, per channel over the whole image,
            "minmax":  (v - min v) / (max v - min v)          with v = (G C R + b) / 255
            "meanstd": clip((v - (m - k s)) / (2 k s), 0, 1)  with m, s = mean, population std of v
            "percentile": clip((v - lo) / (hi - lo), 0, 1)    with lo / hi = the r-th smallest / largest v,
                          r = floor(N k / 100) of the N pixels ("simplest colour balance", Limare et al. 2011)
"""
from __future__ import annotations

import math

import numpy as np

MSRCR_DEFAULTS = dict(
    sigmas=(15.0, 80.0, 250.0),
    weights=(1.0 / 3.0, 1.0 / 3.0, 1.0 / 3.0),
    truncate=4.0,
    alpha=125.0,
    beta=46.0,
    gain=0.6,
    offset=128.0,
    in_scale=255.0,
    in_bias=1.0,
    finish="none",   # "none" | "minmax" | "meanstd" | "percentile" (per image, per channel)
    finish_k=2.0,
)


def gaussian_kernel_1d(sigma: float, truncate: float = 4.0) -> np.ndarray:
    """Sampled, truncated, sum-normalised Gaussian (SURVEY.md 9.1 line 'k_sigma')."""
    r = int(math.ceil(truncate * float(sigma)))
    t = np.arange(-r, r + 1, dtype=np.float64)
    k = np.exp(-(t * t) / (2.0 * float(sigma) ** 2))
    return k / k.sum()


def _blur_axis_direct(a: np.ndarray, k: np.ndarray, axis: int) -> np.ndarray:
    """Definition: shifted-slice FIR with edge clamp along `axis` (float64)."""
    r = (len(k) - 1) // 2
    n = a.shape[axis]
    pad = [(0, 0)] * a.ndim
    pad[axis] = (r, r)
    p = np.pad(a, pad, mode="edge")
    out = np.zeros_like(a)
    sl = [slice(None)] * a.ndim
    for t in range(2 * r + 1):
        sl[axis] = slice(t, t + n)
        out += k[t] * p[tuple(sl)]
    return out


def _blur_axis_fft(a: np.ndarray, k: np.ndarray, axis: int) -> np.ndarray:
    """Same linear operator evaluated by FFT (exact to ~1e-15 relative)."""
    import scipy.fft as sfft

    r = (len(k) - 1) // 2
    n = a.shape[axis]
    pad = [(0, 0)] * a.ndim
    pad[axis] = (r, r)
    p = np.pad(a, pad, mode="edge")
    m = sfft.next_fast_len(n + 4 * r, real=True)
    fa = sfft.rfft(p, m, axis=axis, workers=-1)
    shape = [1] * a.ndim
    shape[axis] = -1
    fk = sfft.rfft(k, m).reshape(shape)
    full = sfft.irfft(fa * fk, m, axis=axis, workers=-1)
    sl = [slice(None)] * a.ndim
    sl[axis] = slice(2 * r, 2 * r + n)  # 'valid' part of conv(p, k) aligned to a
    return full[tuple(sl)]


def blur_edge_clamp(img: np.ndarray, sigma: float, truncate: float = 4.0,
                    mode: str = "direct") -> np.ndarray:
    """Separable Gaussian surround of an (H, W, C) float64 array, edge-clamped."""
    k = gaussian_kernel_1d(sigma, truncate)
    f = _blur_axis_direct if mode == "direct" else _blur_axis_fft
    return f(f(img, k, 1), k, 0)


def msrcr_unclipped(x: np.ndarray, params: dict | None = None, mode: str = "fft") -> np.ndarray:
    """v = (G C R + b) / 255 of one (H, W, 3) image, float64, before any clip or finisher."""
    p = dict(MSRCR_DEFAULTS)
    if params:
        p.update(params)
    x = np.asarray(x)
    assert x.ndim == 3 and x.shape[2] == 3, x.shape
    I = p["in_scale"] * x.astype(np.float64) + p["in_bias"]
    logI = np.log(I)
    R = np.zeros_like(I)
    for s, w in zip(p["sigmas"], p["weights"]):
        S = blur_edge_clamp(I, s, p["truncate"], mode)
        R += w * (logI - np.log(S))
    C = p["beta"] * (np.log(p["alpha"] * I) - np.log(I.sum(axis=2, keepdims=True)))
    return (p["gain"] * C * R + p["offset"]) / 255.0


def msrcr(x: np.ndarray, params: dict | None = None, mode: str = "fft") -> np.ndarray:
    """MSRCR on (H, W, 3) or (N, H, W, 3) float32 in [0, 1] -> float32 in [0, 1]."""
    p = dict(MSRCR_DEFAULTS)
    if params:
        p.update(params)
    x = np.asarray(x)
    if x.ndim == 4:
        return np.stack([msrcr(xi, p, mode) for xi in x])
    out = msrcr_unclipped(x, p, mode)
    if p["finish"] != "none":
        # global-statistics finisher: stretch the UN-clipped value per channel over the whole image
        if p["finish"] == "minmax":
            lo, hi = out.min(axis=(0, 1)), out.max(axis=(0, 1))
        elif p["finish"] == "meanstd":
            m, sd = out.mean(axis=(0, 1)), out.std(axis=(0, 1))   # population standard deviation
            lo, hi = m - p["finish_k"] * sd, m + p["finish_k"] * sd
        elif p["finish"] == "percentile":
            # "simplest colour balance": clip finish_k percent of the pixels at each end (exact order statistics)
            flat = np.sort(out.reshape(-1, 3), axis=0)
            n = flat.shape[0]
            r = int(np.floor(n * float(np.float32(p["finish_k"])) / 100.0))
            r = min(max(r, 0), (n - 1) // 2)
            lo, hi = flat[r], flat[n - 1 - r]
        else:
            raise ValueError(p["finish"])
        out = np.where(hi > lo, (out - lo) / np.where(hi > lo, hi - lo, 1.0), 0.0)
    return np.clip(out, 0.0, 1.0).astype(np.float32)


def msrcr_window(x: np.ndarray, y0: int, y1: int, x0: int, x1: int,
                 params: dict | None = None, mode: str = "fft") -> np.ndarray:
    """Windowed oracle: exact output window [y0,y1) x [x0,x1).

    Crops the input to the window grown by the largest blur radius (clamped to
    the image), runs the unmodified oracle on the crop and returns the window.
    """
    p = dict(MSRCR_DEFAULTS)
    if params:
        p.update(params)
    H, W = x.shape[:2]
    rmax = max(int(math.ceil(p["truncate"] * s)) for s in p["sigmas"])
    cy0, cy1 = max(0, y0 - rmax), min(H, y1 + rmax)
    cx0, cx1 = max(0, x0 - rmax), min(W, x1 + rmax)
    full = msrcr(np.ascontiguousarray(x[cy0:cy1, cx0:cx1]), p, mode)
    return full[y0 - cy0:y1 - cy0, x0 - cx0:x1 - cx0]
