"""Design probe (CPU, float64): how close is the pyramid blur to the oracle's 4-sigma FIR?

Builds the 1-D linear operators of the multi-resolution scheme used by the CUDA
MSRCR path (Gaussian prefilter + decimate by 4 [twice for level 2], low-res
Gaussian, Gaussian-weighted upsample) as dense matrices and compares them with
the oracle's edge-clamped truncated-Gaussian operator: max row L1 error.
"""
import math, sys
import numpy as np

def gk(sigma, r, offs=0.0):
    t = np.arange(-r, r + 1, dtype=np.float64) - offs
    k = np.exp(-t * t / (2 * sigma * sigma))
    return k / k.sum()

def oracle_op(W, sigma, trunc=4.0):
    r = int(math.ceil(trunc * sigma)); k = gk(sigma, r)
    M = np.zeros((W, W))
    for t in range(-r, r + 1):
        idx = np.clip(np.arange(W) + t, 0, W - 1)
        np.add.at(M, (np.arange(W), idx), k[t + r])
    return M

def down_op(n_in, n_out, P_in, P_out, sd, R):
    """out[j'] = sum_t k[t] in[clamp(4(j'-P_out)+P_in+t)]"""
    k = gk(sd, R); M = np.zeros((n_out, n_in))
    for jo in range(n_out):
        c = 4 * (jo - P_out) + P_in
        idx = np.clip(c + np.arange(-R, R + 1), 0, n_in - 1)
        np.add.at(M[jo], idx, k)
    return M

def blur_op(n, sigma, r):
    k = gk(sigma, r); M = np.zeros((n, n))
    for t in range(-r, r + 1):
        idx = np.clip(np.arange(n) + t, 0, n - 1)
        np.add.at(M, (np.arange(n), idx), k[t + r])
    return M

def up_op(W, n, s, P, su, T):
    """out[X] = sum_{j=j0-T..j0+T} w[j] L[clamp j], u = X/s + P, j0 = round(u)"""
    M = np.zeros((W, n))
    for X in range(W):
        u = X / s + P; j0 = int(math.floor(u + 0.5))
        j = np.arange(j0 - T, j0 + T + 1)
        w = np.exp(-(j - u) ** 2 / (2 * su * su)); w /= w.sum()
        np.add.at(M[X], np.clip(j, 0, n - 1), w)
    return M

def kvar(k):
    t = np.arange(len(k)) - (len(k) - 1) / 2
    return (k * t * t).sum()

def build(W, sigma, level, sd=4.0, Rd=16, su=1.0, T=4, rl_mult=4.0, match_var=True, trunc=4.0):
    s = 4 ** level
    P1 = Rd // 4 + (1 if Rd % 4 else 0); P1 = max(P1, T)
    n1 = P1 + (W - 1 + Rd + 3) // 4 + 1 + T
    D = down_op(W, n1, 0, P1, sd, Rd)
    var = kvar(gk(sd, Rd)); n, P = n1, P1
    if level == 2:
        ext = Rd + 4 * Rd
        P2 = max((ext + 15) // 16, T)
        n2 = P2 + (W - 1 + ext + 15) // 16 + 1 + T
        D = down_op(n1, n2, P1, P2, sd, Rd) @ D
        var += 16 * kvar(gk(sd, Rd)); n, P = n2, P2
    # phase-averaged variance of the upsampler (full-res px^2)
    vs = []
    for ph in range(s):
        u = ph / s; j0 = math.floor(u + 0.5); j = np.arange(j0 - T, j0 + T + 1)
        w = np.exp(-(j - u) ** 2 / (2 * su * su)); w /= w.sum()
        m = (w * (j - u)).sum(); vs.append(((w * (j - u) ** 2).sum() - 0*m*m) * s * s)
    var_up = float(np.mean(vs))
    r = int(math.ceil(trunc * sigma))
    target = kvar(gk(sigma, r)) if match_var else sigma ** 2
    sl = math.sqrt(max(target - var - var_up, 1e-9)) / s
    # iterate once: truncated low-res kernel has smaller variance than sl^2
    rl = int(math.ceil(rl_mult * sl))
    for _ in range(20):
        v = kvar(gk(sl, rl)) * s * s
        sl *= math.sqrt((target - var - var_up) / v)
    B = blur_op(n, sl, rl)
    U = up_op(W, n, s, P, su, T)
    return U @ B @ D, dict(sl=sl, rl=rl, n=n, P=P, var_pre=var, var_up=var_up)

if __name__ == "__main__":
    np.set_printoptions(linewidth=200)
    for sigma, level, W in ((15, 1, 1024), (80, 2, 2048), (250, 2, 4096), (250, 2, 600)):
        Mo = oracle_op(W, sigma)
        for kw in (dict(), dict(rl_mult=3.5), dict(rl_mult=4.5), dict(sd=3.0, Rd=12), dict(T=3), dict(su=0.8, T=3),dict(sd=3.0, Rd=12,su=0.8,T=3)):
            M, info = build(W, sigma, level, **kw)
            e = np.abs(M - Mo).sum(1)
            print(f"sigma={sigma} L{level} W={W} {kw}: L1 max {e.max():.3e} mean {e.mean():.3e} rowsum-1 {np.abs(M.sum(1)-1).max():.1e} sl={info['sl']:.3f} rl={info['rl']}")
