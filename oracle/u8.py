"""8-bit image convention of the oracle.  Self-written oracle -- test infrastructure,
not code from CILAB-IITM/Low_Light_Image_Enhancement (which holds no code: /root/reference/README.md:1-2).

A byte b means the value float32(b) / 255; a result v in [0, 1] is stored as rint(255 * v) (round half to even).
16-bit images follow the same convention with 65535.
"""
from __future__ import annotations

import numpy as np

from .msrcr import msrcr
from .zero_dce import zero_dce


def from_u8(x_u8: np.ndarray) -> np.ndarray:
    assert x_u8.dtype == np.uint8
    return (x_u8.astype(np.float32) / np.float32(255.0)).astype(np.float32)


def to_u8(v: np.ndarray) -> np.ndarray:
    return np.rint(np.clip(np.asarray(v, dtype=np.float64), 0.0, 1.0) * 255.0).astype(np.uint8)


def msrcr_u8(x_u8: np.ndarray, params: dict | None = None, mode: str = "fft"):
    """Returns (uint8 result, the un-rounded float64 result in 0..255 units)."""
    y = msrcr(from_u8(x_u8), params, mode=mode)
    return to_u8(y), np.clip(np.asarray(y, dtype=np.float64), 0, 1) * 255.0


def zero_dce_u8(x_u8: np.ndarray, ws, **kw):
    y = zero_dce(from_u8(x_u8), ws, **kw)
    return to_u8(y), np.clip(np.asarray(y, dtype=np.float64), 0, 1) * 255.0


def from_u16(x_u16: np.ndarray) -> np.ndarray:
    assert x_u16.dtype == np.uint16
    return (x_u16.astype(np.float32) / np.float32(65535.0)).astype(np.float32)


def to_u16(v: np.ndarray) -> np.ndarray:
    return np.rint(np.clip(np.asarray(v, dtype=np.float64), 0.0, 1.0) * 65535.0).astype(np.uint16)


def msrcr_u16(x_u16: np.ndarray, params: dict | None = None, mode: str = "fft"):
    """-> (uint16 result, the un-rounded 65535 * value as float64, for tie-aware comparisons)."""
    y = msrcr(from_u16(x_u16), params, mode=mode)
    return to_u16(y), np.clip(np.asarray(y, dtype=np.float64), 0, 1) * 65535.0


def zero_dce_u16(x_u16: np.ndarray, ws, **kw):
    y = zero_dce(from_u16(x_u16), ws, **kw)
    return to_u16(y), np.clip(np.asarray(y, dtype=np.float64), 0, 1) * 65535.0
