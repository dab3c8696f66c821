"""NumPy emulation of the bf16 Zero-DCE path as SURVEY.md 9.2 defines it (test helper, CPU):
the input image, every inter-layer activation and all seven layers' weights rounded to bf16 (round-to-nearest-even), fp32-or-better accumulation, bias/ReLU/tanh/
curves in full precision.  Used to localise errors layer by layer; the acceptance bar itself
is PSNR >= 50 dB against the un-rounded fp64 oracle."""
import numpy as np

import oracle
from oracle.zero_dce import _conv3x3


def bf16_round(a: np.ndarray) -> np.ndarray:
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32)


def dce_activations_bf16(x_hwc: np.ndarray, ws):
    """Returns [x1..x6] (each (32,H,W) float32, bf16-representable) and A (24,H,W) float64."""
    relu = lambda a: np.maximum(a, 0)
    q = lambda w, b: (bf16_round(w).astype(np.float64), b.astype(np.float64))
    x = bf16_round(x_hwc.transpose(2, 0, 1)).astype(np.float64)
    x1 = bf16_round(relu(_conv3x3(x, *q(*ws[0]))))
    f = lambda z, i: bf16_round(relu(_conv3x3(z.astype(np.float64), *q(*ws[i]))))
    x2 = f(x1, 1); x3 = f(x2, 2); x4 = f(x3, 3)
    x5 = f(np.concatenate([x3, x4]), 4)
    x6 = f(np.concatenate([x2, x5]), 5)
    A = np.tanh(_conv3x3(np.concatenate([x1, x6]).astype(np.float64), *q(*ws[6])))
    return [x1, x2, x3, x4, x5, x6], A
