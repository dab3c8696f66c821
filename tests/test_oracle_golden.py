"""The oracle against the committed golden vectors (tests/golden/make_golden.py)."""
import os

import numpy as np

import oracle

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_msrcr_golden_default():
    d = np.load(os.path.join(G, "msrcr_96x128.npz"))
    assert np.array_equal(oracle.msrcr(d["x"], mode="direct"), d["y"])
    assert np.abs(oracle.msrcr(d["x"], mode="fft") - d["y"]).max() < 1e-6
    assert np.array_equal(oracle.synth_lowlight(96, 128, seed=11), d["x"])


def test_msrcr_golden_custom_scales():
    d = np.load(os.path.join(G, "msrcr_80x112_custom.npz"))
    p = dict(sigmas=tuple(d["sigmas"]), weights=tuple(d["weights"]), gain=float(d["gain"]), offset=float(d["offset"]))
    assert np.array_equal(oracle.msrcr(d["x"], p, mode="direct"), d["y"])


def test_zero_dce_golden():
    d = np.load(os.path.join(G, "zero_dce_48x64.npz"))
    ws = oracle.unpack_weights(d["weights"])
    y, A = oracle.zero_dce(d["x"], ws, return_maps=True)
    assert np.abs(y - d["y"]).max() < 1e-7 and np.abs(A - d["maps"]).max() < 1e-7
    assert np.array_equal(d["weights"], oracle.pack_weights(oracle.init_weights(0)))
