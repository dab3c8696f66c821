"""Design probe: tune (sigma_low, r_low, end-tap fraction) of the low-res stage to minimise the
row-L1 distance between the pyramid's effective kernel and the oracle's truncated Gaussian."""
import sys, math, numpy as np
sys.path.insert(0, '.')
from tools.pyramid_probe import gk, oracle_op, down_op, up_op, kvar
from scipy.optimize import minimize

def setup(W, level, sd=4.0, Rd=16, su=1.0, T=4):
    s = 4 ** level
    P1 = max((Rd + 3) // 4, T); n1 = P1 + (W - 1 + Rd + 3) // 4 + 1 + T
    D = down_op(W, n1, 0, P1, sd, Rd); n, P = n1, P1
    if level == 2:
        ext = Rd + 4 * Rd; P2 = max((ext + 15) // 16, T); n2 = P2 + (W - 1 + ext + 15) // 16 + 1 + T
        D = down_op(n1, n2, P1, P2, sd, Rd) @ D; n, P = n2, P2
    U = up_op(W, n, s, P, su, T)
    return D, U, n, s

def lowk(sl, rl, f):
    t = np.arange(-rl, rl + 1, dtype=np.float64); k = np.exp(-t * t / (2 * sl * sl)); k[0] *= f; k[-1] *= f
    return k / k.sum()

def eff_rows(D, U, n, rows, k):
    rl = (len(k) - 1) // 2
    B = np.zeros((n, n))
    for t in range(-rl, rl + 1):
        idx = np.clip(np.arange(n) + t, 0, n - 1); np.add.at(B, (np.arange(n), idx), k[t + rl])
    return (U[rows] @ B) @ D

for sigma, level, W in ((15, 1, 768), (80, 2, 2048), (250, 2, 4608)):
    D, U, n, s = setup(W, level)
    Mo = oracle_op(W, sigma)
    rows = np.arange(W // 2, W // 2 + s)
    best = None
    sl0 = math.sqrt(sigma ** 2 - (16 if level == 1 else 272) - s * s) / s
    for rl in range(int(3.5 * sl0), int(4.6 * sl0) + 2):
        def obj(p):
            k = lowk(p[0], rl, min(max(p[1], 0.0), 1.0))
            return np.abs(eff_rows(D, U, n, rows, k) - Mo[rows]).sum(1).mean()
        for f0 in (0.3, 0.7, 1.0):
            r = minimize(obj, [sl0, f0], method='Nelder-Mead', options=dict(xatol=1e-5, fatol=1e-9))
            if best is None or r.fun < best[0]:
                best = (r.fun, rl, r.x[0], min(max(r.x[1], 0), 1))
        print(sigma, 'rl', rl, 'best so far', best)
    print('RESULT', sigma, level, best)
