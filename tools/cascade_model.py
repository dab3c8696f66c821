"""Float64 twin of the MSRCR cascade (csrc/msrcr_kernels.cuh), for design work and CPU tests of the planner.

Builds the 1-D linear operator pixels -> surround of every stage of the GPU path from the index formulas of the
kernels (prefilter + decimate to level C [and B], the planner's low-resolution taps, the Gaussian-weighted B -> C
upsampler, the final Lagrange step), composes them into one dense matrix per scale, and evaluates the full MSRCR
formula with them.  Exact arithmetic of the same operators: what remains against the oracle is the cascade's
kernel error, without any fp32 effect.  Needs libllie.so only for the planner (no GPU).
"""
from __future__ import annotations

import functools
import math
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

kPC, kPB, kRD, kTU = 4, 5, 16, 4


def _gauss(sigma, r):
    t = np.arange(-r, r + 1, dtype=np.float64)
    k = np.exp(-t * t / (2 * sigma * sigma))
    return k / k.sum()


def _lagrange(T, frac):
    w = np.ones(T)
    for i in range(T):
        oi = i - (T // 2 - 1)
        for m in range(T):
            if m != i:
                om = m - (T // 2 - 1)
                w[i] *= (frac - om) / (oi - om)
    return w


def level_c_size(n):
    return kPC + (n + 18) // 4 + 1


def level_b_size(n):
    return kPB + (n + 94) // 16 + 1


def op_down_c(n):
    """C[j] = sum_t g[t] I[clamp(4 (j - PC) + t - 16)]"""
    g, nc = _gauss(4.0, kRD), level_c_size(n)
    M = np.zeros((nc, n))
    for j in range(nc):
        idx = np.clip(4 * (j - kPC) + np.arange(-kRD, kRD + 1), 0, n - 1)
        np.add.at(M[j], idx, g)
    return M


def op_down_b(n):
    """B[j] = sum_t g[t] C[clamp(4 (j - PB) + PC + t - 16)]"""
    g, nc, nb = _gauss(4.0, kRD), level_c_size(n), level_b_size(n)
    M = np.zeros((nb, nc))
    for j in range(nb):
        idx = np.clip(4 * (j - kPB) + kPC + np.arange(-kRD, kRD + 1), 0, nc - 1)
        np.add.at(M[j], idx, g)
    return M


def op_fir(m, taps):
    r = (len(taps) - 1) // 2
    M = np.zeros((m, m))
    for j in range(m):
        np.add.at(M[j], np.clip(j + np.arange(-r, r + 1), 0, m - 1), taps)
    return M


def op_up_bc(n):
    nc, nb = level_c_size(n), level_b_size(n)
    M = np.zeros((nc, nb))
    for jc in range(nc):
        d = jc - kPC + 4
        ph = d & 3
        j0 = (d >> 2) - 1 + kPB + (1 if ph >= 2 else 0)
        shift = (1.0 if ph >= 2 else 0.0) - ph / 4.0
        dd = shift - kTU + np.arange(2 * kTU + 1)
        w = np.exp(-dd * dd / 2.0)
        np.add.at(M[jc], np.clip(j0 - kTU + np.arange(2 * kTU + 1), 0, nb - 1), w / w.sum())
    return M


def op_final(n, T):
    nc = level_c_size(n)
    M = np.zeros((n, nc))
    for y in range(n):
        j0, ph = y // 4 + kPC, y & 3
        if ph == 0:
            M[y, min(j0, nc - 1)] += 1.0
        else:
            w = _lagrange(T, ph / 4.0)
            np.add.at(M[y], np.minimum(j0 - (T // 2 - 1) + np.arange(T), nc - 1), w)
    return M


@functools.lru_cache(maxsize=32)
def surround_operator(n: int, sigma: float, truncate: float = 4.0) -> np.ndarray:
    """(n, n) matrix of the cascade's 1-D surround operator for an axis of n pixels."""
    import low_light_image_enhancement_b200 as L
    q = L.plan_query(sigma, truncate)
    taps = q["taps"].astype(np.float64)
    C = op_down_c(n)
    if q["stride"] == 4:
        return op_final(n, 6) @ op_fir(level_c_size(n), taps) @ C
    return op_final(n, 4) @ op_up_bc(n) @ op_fir(level_b_size(n), taps) @ op_down_b(n) @ C


def oracle_operator(n: int, sigma: float, truncate: float = 4.0) -> np.ndarray:
    r = int(math.ceil(truncate * sigma))
    k = _gauss(sigma, r)
    M = np.zeros((n, n))
    for y in range(n):
        np.add.at(M[y], np.clip(y + np.arange(-r, r + 1), 0, n - 1), k)
    return M


def msrcr_cascade(x: np.ndarray, params: dict | None = None) -> np.ndarray:
    """MSRCR of one (H, W, 3) image with the cascade's operators in float64 (un-finished: plain clip)."""
    import oracle
    p = dict(oracle.MSRCR_DEFAULTS)
    if params:
        p.update(params)
    I = p["in_scale"] * x.astype(np.float64) + p["in_bias"]
    H, W = I.shape[:2]
    logI, R = np.log(I), np.zeros_like(I)
    for s, wt in zip(p["sigmas"], p["weights"]):
        My, Mx = surround_operator(H, s, p["truncate"]), surround_operator(W, s, p["truncate"])
        S = np.einsum("xw,ywc->yxc", Mx, np.einsum("yh,hwc->ywc", My, I))
        R += wt * (logI - np.log(S))
    Cc = p["beta"] * (np.log(p["alpha"] * I) - np.log(I.sum(axis=2, keepdims=True)))
    return np.clip((p["gain"] * Cc * R + p["offset"]) / 255.0, 0.0, 1.0)


if __name__ == "__main__":
    import oracle
    h, w = (int(v) for v in (sys.argv[1:3] if len(sys.argv) > 2 else (480, 720)))
    for name, x in oracle.hard_cases(h, w).items():
        ref = oracle.msrcr(x).astype(np.float64)
        out = msrcr_cascade(x)
        d = np.abs(out - ref)
        iy, ix, ic = np.unravel_index(d.argmax(), d.shape)
        print(f"{name:24s} [{oracle.HARD_CASE_CLASS[name]:4s}] max-abs {d.max():.2e} at (y={iy}, x={ix}, c={ic})  mean {d.mean():.2e}")
    x = oracle.synth_lowlight(h, w, 3)
    print(f"{'synth':24s} max-abs {np.abs(msrcr_cascade(x) - oracle.msrcr(x)).max():.2e}")
