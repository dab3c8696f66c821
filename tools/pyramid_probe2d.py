"""Design probe: full MSRCR output error of the pyramid scheme (float64 operators) vs oracle."""
import sys, math, numpy as np
sys.path.insert(0, '.')
import oracle
from tools.pyramid_probe import build
def msrcr_pyr(x, kw):
    p = oracle.MSRCR_DEFAULTS
    I = 255.0 * x.astype(np.float64) + 1.0
    H, W = I.shape[:2]
    logI = np.log(I); R = np.zeros_like(I)
    for s, lvl in ((15, 1), (80, 2), (250, 2)):
        My, _ = build(H, s, lvl, **kw); Mx, _ = build(W, s, lvl, **kw)
        S = np.einsum('yh,hwc->ywc', My, I); S = np.einsum('xw,ywc->yxc', Mx, S)
        R += (logI - np.log(S)) / 3
    C = p['beta'] * (np.log(p['alpha'] * I) - np.log(I.sum(2, keepdims=True)))
    return np.clip((p['gain'] * C * R + p['offset']) / 255, 0, 1)
for (h, w, seed, scale) in ((400, 600, 0, 1), (720, 1280, 3, 1), (720, 1280, 4, 6.0), (1080, 1920, 5, 1)):
    x = np.clip(oracle.synth_lowlight(h, w, seed) * scale, 0, 1).astype(np.float32)
    ref = oracle.msrcr(x).astype(np.float64)
    for kw in (dict(), dict(sd=3.0, Rd=12, su=0.8, T=3), dict(rl_mult=3.8)):
        out = msrcr_pyr(x, kw)
        d = np.abs(out - ref)
        print(h, w, seed, scale, kw, 'max', d.max(), 'mean', d.mean())
