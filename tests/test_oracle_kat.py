"""Closed-form known-answer tests of the oracle.  The reference ships no tests or
golden vectors (/root/reference/README.md:1-2); these pin the oracle's arithmetic instead."""
import math

import numpy as np
import pytest

import oracle


def test_msrcr_constant_image_gives_offset():
    # every S_sigma equals I on a constant image => R = 0 => out = offset / 255, for any sigma
    for val in (0.0, 0.2, 1.0):
        x = np.full((40, 56, 3), val, np.float32)
        y = oracle.msrcr(x, mode="direct")
        assert np.abs(y - 128.0 / 255.0).max() < 1e-7
        y = oracle.msrcr(x, dict(sigmas=(3.0, 7.0), weights=(0.5, 0.5), offset=100.0), mode="fft")
        assert np.abs(y - 100.0 / 255.0).max() < 1e-7


def test_msrcr_grey_image_colour_restoration_is_constant():
    # R=G=B => C = beta ln(alpha/3) everywhere, channels identical
    g = oracle.synth_lowlight(32, 48, 3)[:, :, :1]
    x = np.repeat(g, 3, axis=2)
    p = oracle.MSRCR_DEFAULTS
    I = 255.0 * x.astype(np.float64) + 1
    R = sum(w * (np.log(I) - np.log(oracle.blur_edge_clamp(I, s))) for s, w in zip(p["sigmas"], p["weights"]))
    expect = np.clip((p["gain"] * p["beta"] * math.log(p["alpha"] / 3.0) * R + p["offset"]) / 255.0, 0, 1)
    y = oracle.msrcr(x)
    assert np.abs(y - expect).max() < 1e-6
    assert np.array_equal(y[:, :, 0], y[:, :, 1]) and np.array_equal(y[:, :, 1], y[:, :, 2])


def test_blur_impulse_is_outer_product_of_the_kernel():
    sigma, n = 3.0, 41
    k = oracle.gaussian_kernel_1d(sigma)
    r = (len(k) - 1) // 2
    assert r == 12 and abs(k.sum() - 1) < 1e-15 and np.allclose(k, k[::-1])
    img = np.zeros((n, n, 1))
    img[n // 2, n // 2, 0] = 1.0
    for mode in ("direct", "fft"):
        b = oracle.blur_edge_clamp(img, sigma, mode=mode)[:, :, 0]
        c = n // 2
        assert np.abs(b[c - r:c + r + 1, c - r:c + r + 1] - np.outer(k, k)).max() < 1e-15


def test_blur_radius_larger_than_image_is_well_defined():
    # config 0: r = 1000 on 600x400; clamp index = clip(i, 0, n-1)
    img = np.arange(12, dtype=np.float64).reshape(3, 4, 1)
    b = oracle.blur_edge_clamp(img, 250.0)
    assert b.shape == img.shape and b.min() >= img.min() and b.max() <= img.max()
    # rows far beyond the kernel centre are weighted like the edges: result is nearly the edge mean
    assert abs(b.mean() - 5.5) < 1e-9


def test_blur_preserves_constants_and_bounds():
    rng = np.random.default_rng(0)
    img = rng.uniform(1, 256, size=(30, 37, 3))
    for s in (2.0, 15.0):
        b = oracle.blur_edge_clamp(img, s, mode="fft")
        assert b.min() >= img.min() - 1e-9 and b.max() <= img.max() + 1e-9
    c = np.full((9, 9, 2), 7.25)
    assert np.abs(oracle.blur_edge_clamp(c, 80.0) - 7.25).max() < 1e-12


def test_msrcr_is_channel_permutation_equivariant():
    x = oracle.synth_lowlight(24, 40, 5)
    y = oracle.msrcr(x)
    perm = [2, 0, 1]
    assert np.array_equal(oracle.msrcr(np.ascontiguousarray(x[:, :, perm])), y[:, :, perm])


def _zero_weights(last_bias=None):
    ws = [(np.zeros((co, ci, 3, 3), np.float32), np.zeros(co, np.float32)) for co, ci in oracle.DCE_LAYERS]
    if last_bias is not None:
        ws[-1] = (ws[-1][0], np.asarray(last_bias, np.float32))
    return ws


def test_dce_zero_weights_is_identity():
    x = oracle.synth_lowlight(16, 20, 1)
    y, A = oracle.zero_dce(x, _zero_weights(), return_maps=True)
    assert np.array_equal(y, x) and np.all(A == 0)


def test_dce_constant_alpha_closed_form():
    b = np.linspace(-0.9, 0.9, 24)
    x = oracle.synth_lowlight(8, 12, 2) * 5
    y, A = oracle.zero_dce(x, _zero_weights(b), return_maps=True)
    assert np.allclose(A, np.tanh(b).astype(np.float32)[None, None, :], atol=1e-7)
    e = x.astype(np.float64)
    a = np.tanh(b.astype(np.float32).astype(np.float64))
    for n in range(8):
        e = e + a[3 * n:3 * n + 3] * e * (1 - e)
    assert np.abs(y - e).max() < 1e-7


def test_curve_alpha_one_and_fixed_points():
    x = np.linspace(0, 1, 33, dtype=np.float64).reshape(1, 1, 33).repeat(3, 0)
    A = np.ones((24, 1, 33))
    assert np.allclose(oracle.apply_curves(x, A, iters=1), 2 * x - x * x)
    ends = np.array([0.0, 1.0]).reshape(1, 1, 2).repeat(3, 0)
    rng = np.random.default_rng(0)
    assert np.array_equal(oracle.apply_curves(ends, rng.uniform(-1, 1, (24, 1, 2))), ends)


def test_curves_stay_in_unit_interval():
    rng = np.random.default_rng(1)
    x = rng.uniform(0, 1, (3, 16, 16))
    A = rng.uniform(-1, 1, (24, 16, 16))
    y = oracle.apply_curves(x, A)
    assert y.min() >= 0 and y.max() <= 1


def test_dce_param_count_and_pack_roundtrip():
    assert oracle.DCE_NUM_PARAMS == 79416
    ws = oracle.init_weights(3)
    blob = oracle.pack_weights(ws)
    for (w, b), (w2, b2) in zip(ws, oracle.unpack_weights(blob)):
        assert np.array_equal(w, w2) and np.array_equal(b, b2)


def test_dce_batch_elements_are_independent():
    ws = oracle.init_weights(0)
    xs = np.stack([oracle.synth_lowlight(20, 24, s) for s in (1, 2)])
    yb = oracle.zero_dce(xs, ws)
    assert np.array_equal(yb[0], oracle.zero_dce(xs[0], ws)) and np.array_equal(yb[1], oracle.zero_dce(xs[1], ws))


def test_synth_lowlight_is_dark_and_seeded():
    a, b = oracle.synth_lowlight(64, 96, 7), oracle.synth_lowlight(64, 96, 7)
    assert np.array_equal(a, b) and a.dtype == np.float32 and a.shape == (64, 96, 3)
    assert 0 <= a.min() and a.max() < 0.2 and a.mean() < 0.06
    assert not np.array_equal(a, oracle.synth_lowlight(64, 96, 8))


def test_u8_convention_round_trips():
    """byte b <-> b / 255, rint on the way back (half to even)."""
    b = np.arange(256, dtype=np.uint8)
    assert np.array_equal(oracle.to_u8(oracle.from_u8(b)), b)
    assert oracle.to_u8(np.array([0.5 / 255, 1.5 / 255, -1.0, 2.0, 0.4999 / 255])).tolist() == [0, 2, 0, 255, 0]


def test_u16_convention_kat():
    b = np.array([0, 1, 32767, 65535], np.uint16)
    assert np.array_equal(oracle.to_u16(oracle.from_u16(b)), b)
    assert oracle.to_u16(np.array([0.5 / 65535, 1.5 / 65535, -1.0, 2.0])).tolist() == [0, 2, 0, 65535]
    x8 = np.clip(np.rint(oracle.synth_lowlight(48, 64, 3) * 6 * 255), 0, 255).astype(np.uint8)
    y16, _ = oracle.msrcr_u16(x8.astype(np.uint16) * 257)       # 257 b / 65535 == b / 255: the same image
    y8, y8f = oracle.msrcr_u8(x8)
    assert np.abs(y16.astype(np.float64) / 257.0 - y8f).max() <= 0.01


def test_msrcr_finishers_kat():
    """min/max stretch reaches exactly 0 and 1 in every channel; mean/std stretch centres the
    un-clipped value at 0.5 and is invariant to gain and offset (an affine map of v cancels in the stretch)."""
    x = oracle.synth_lowlight(64, 80, 5)
    y = oracle.msrcr(x, dict(finish="minmax"))
    assert np.all(y.min(axis=(0, 1)) == 0.0) and np.all(y.max(axis=(0, 1)) == 1.0)
    a = oracle.msrcr(x, dict(finish="meanstd", finish_k=3.0))
    b = oracle.msrcr(x, dict(finish="meanstd", finish_k=3.0, gain=0.3, offset=60.0))
    assert np.abs(a - b).max() <= 1e-6
    assert np.abs(a.mean(axis=(0, 1)) - 0.5).max() <= 0.05     # clipping of the tails moves the mean only slightly
    # percentile stretch: k = 0 is the min/max stretch; k percent of the pixels saturate at each end
    assert np.array_equal(oracle.msrcr(x, dict(finish="percentile", finish_k=0.0)), y)
    q = oracle.msrcr(x, dict(finish="percentile", finish_k=5.0))
    n, r = 64 * 80, int(64 * 80 * 5.0 / 100.0)
    for c in range(3):
        assert (q[..., c] == 0.0).sum() >= r + 1 and (q[..., c] == 1.0).sum() >= r + 1
        assert (q[..., c] == 0.0).sum() <= r + 8 and (q[..., c] == 1.0).sum() <= r + 8    # ties only
    assert np.abs(q - oracle.msrcr(x, dict(finish="percentile", finish_k=5.0, gain=0.3, offset=60.0))).max() <= 1e-6
