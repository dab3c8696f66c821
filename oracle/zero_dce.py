"""NumPy oracle for Zero-DCE (BASELINE.json north_star (b)).

Self-written oracle -- not code from CILAB-IITM/Low_Light_Image_Enhancement
(the reference holds no code: /root/reference/README.md:1-2).  Test
infrastructure only.

    x1 = relu(conv(x, W1)), x2..x4 = relu(conv(prev)), x5 = relu(conv(cat(x3,x4))),
    x6 = relu(conv(cat(x2,x5))), A = tanh(conv(cat(x1,x6)))   (3x3, stride 1, zero pad 1)
    y0 = x; y_{n+1} = y_n + A[3n:3n+3] * y_n * (1 - y_n), n = 0..7
"""
from __future__ import annotations

import numpy as np

# (Cout, Cin) per layer; W stored OIHW (Cout, Cin, 3, 3), then bias (Cout,)
DCE_LAYERS = ((32, 3), (32, 32), (32, 32), (32, 32), (32, 64), (32, 64), (24, 64))
DCE_NUM_PARAMS = sum(co * ci * 9 + co for co, ci in DCE_LAYERS)  # 79 416


def init_weights(seed: int = 0, last_gain: float = 1.0):
    """He-normal W ~ N(0, 2/(9 Cin)), b ~ U(-0.05, 0.05), drawn W1,b1,...,W7,b7."""
    rng = np.random.default_rng(seed)
    ws = []
    for co, ci in DCE_LAYERS:
        w = rng.normal(0.0, np.sqrt(2.0 / (9.0 * ci)), size=(co, ci, 3, 3)).astype(np.float32)
        b = rng.uniform(-0.05, 0.05, size=(co,)).astype(np.float32)
        ws.append((w, b))
    if last_gain != 1.0:
        w, b = ws[-1]
        ws[-1] = ((w * np.float32(last_gain)).astype(np.float32), b)
    return ws


def pack_weights(ws) -> np.ndarray:
    """Flat float32 blob W1 b1 ... W7 b7 (79 416 floats) shared with llie_dce_set_weights."""
    blob = np.concatenate([np.concatenate([w.ravel(), b.ravel()]) for w, b in ws]).astype(np.float32)
    assert blob.size == DCE_NUM_PARAMS
    return blob


def unpack_weights(blob: np.ndarray):
    blob = np.asarray(blob, dtype=np.float32).ravel()
    assert blob.size == DCE_NUM_PARAMS, blob.size
    ws, o = [], 0
    for co, ci in DCE_LAYERS:
        w = blob[o:o + co * ci * 9].reshape(co, ci, 3, 3); o += co * ci * 9
        b = blob[o:o + co]; o += co
        ws.append((w, b))
    return ws


def _conv3x3(z: np.ndarray, w: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Cross-correlation, zero pad 1: z (Cin,H,W) -> (Cout,H,W), dtype of z."""
    ci, h, wd = z.shape
    zp = np.zeros((ci, h + 2, wd + 2), dtype=z.dtype)
    zp[:, 1:-1, 1:-1] = z
    out = np.empty((w.shape[0], h * wd), dtype=z.dtype)
    out[:] = b.astype(z.dtype)[:, None]
    wt = w.astype(z.dtype)
    for dy in range(3):
        for dx in range(3):
            out += wt[:, :, dy, dx] @ zp[:, dy:dy + h, dx:dx + wd].reshape(ci, h * wd)
    return out.reshape(w.shape[0], h, wd)


def dce_net(x_chw: np.ndarray, ws, dtype=np.float64) -> np.ndarray:
    """(3,H,W) in [0,1] -> curve maps A (24,H,W) in (-1,1)."""
    relu = lambda a: np.maximum(a, 0)
    x = x_chw.astype(dtype)
    x1 = relu(_conv3x3(x, *ws[0]))
    x2 = relu(_conv3x3(x1, *ws[1]))
    x3 = relu(_conv3x3(x2, *ws[2]))
    x4 = relu(_conv3x3(x3, *ws[3]))
    x5 = relu(_conv3x3(np.concatenate([x3, x4]), *ws[4]))
    x6 = relu(_conv3x3(np.concatenate([x2, x5]), *ws[5]))
    return np.tanh(_conv3x3(np.concatenate([x1, x6]), *ws[6]))


def apply_curves(x_chw: np.ndarray, A: np.ndarray, iters: int = 8, sign: int = 1) -> np.ndarray:
    y = x_chw.astype(A.dtype)
    for n in range(iters):
        y = y + sign * A[3 * n:3 * n + 3] * y * (1.0 - y)
    return y


def zero_dce(x: np.ndarray, ws, iters: int = 8, dtype=np.float64, return_maps: bool = False,
             band: int = 0):
    """Zero-DCE on (H,W,3) / (N,H,W,3) float32 -> float32 same layout.

    `band` > 0 evaluates in row bands of that height with a 7-row halo (exact:
    the receptive field of the 7 stacked 3x3 convs is 15x15) to bound memory.
    """
    x = np.asarray(x)
    if x.ndim == 4:
        outs = [zero_dce(xi, ws, iters, dtype, return_maps, band) for xi in x]
        if return_maps:
            return np.stack([o[0] for o in outs]), np.stack([o[1] for o in outs])
        return np.stack(outs)
    assert x.ndim == 3 and x.shape[2] == 3, x.shape
    H = x.shape[0]
    xc = np.ascontiguousarray(x.transpose(2, 0, 1))
    if band and band < H:
        ys, As = [], []
        for y0 in range(0, H, band):
            y1 = min(H, y0 + band)
            c0, c1 = max(0, y0 - 7), min(H, y1 + 7)
            # zero padding of the crop equals the net's zero padding only at true borders;
            # elsewhere the 7-row halo shields the kept rows from it.
            A = dce_net(xc[:, c0:c1], ws, dtype)[:, y0 - c0:y1 - c0]
            ys.append(apply_curves(xc[:, y0:y1], A, iters))
            As.append(A)
        y = np.concatenate(ys, axis=1)
        A = np.concatenate(As, axis=1)
    else:
        A = dce_net(xc, ws, dtype)
        y = apply_curves(xc, A, iters)
    y = y.transpose(1, 2, 0).astype(np.float32)
    if return_maps:
        return y, A.transpose(1, 2, 0).astype(np.float32)
    return y
