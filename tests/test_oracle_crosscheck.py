"""The oracle against independent implementations ('parity unpinned by the
reference' is compensated by cross-library agreement): SciPy ndimage, OpenCV, torch CPU."""
import numpy as np
import pytest

import oracle


def _img(h, w, seed):
    return 255.0 * oracle.synth_lowlight(h, w, seed).astype(np.float64) + 1.0


def test_direct_and_fft_modes_agree():
    I = _img(60, 90, 0)
    for s in (2.5, 15.0, 80.0):
        a = oracle.blur_edge_clamp(I, s, mode="direct")
        b = oracle.blur_edge_clamp(I, s, mode="fft")
        assert np.abs(a - b).max() / np.abs(a).max() < 1e-13


def test_blur_matches_scipy_gaussian_filter():
    ndi = pytest.importorskip("scipy.ndimage")
    I = _img(70, 100, 1)
    for s in (1.5, 6.0, 15.0):
        ref = np.stack([ndi.gaussian_filter(I[:, :, c], s, mode="nearest", truncate=4.0) for c in range(3)], axis=2)
        assert np.abs(oracle.blur_edge_clamp(I, s) - ref).max() / ref.max() < 1e-12


def test_blur_matches_opencv_border_replicate():
    cv2 = pytest.importorskip("cv2")
    I = _img(64, 80, 2)
    for s in (2.0, 7.0):
        r = int(np.ceil(4 * s))
        ref = cv2.GaussianBlur(I, (2 * r + 1, 2 * r + 1), sigmaX=s, sigmaY=s, borderType=cv2.BORDER_REPLICATE)
        assert np.abs(oracle.blur_edge_clamp(I, s) - ref).max() / ref.max() < 1e-12


def test_windowed_oracle_is_bit_identical_to_full_image():
    x = oracle.synth_lowlight(300, 380, 4)
    p = dict(sigmas=(3.0, 12.0, 30.0))  # r_max = 120: windows below exercise cut and uncut sides
    full = oracle.msrcr(x, p)
    for (y0, x0) in ((0, 0), (0, 316), (236, 0), (236, 316), (130, 150), (0, 170), (100, 0)):
        win = oracle.msrcr_window(x, y0, y0 + 64, x0, x0 + 64, p)
        assert np.array_equal(win, full[y0:y0 + 64, x0:x0 + 64])


def test_dce_net_matches_torch_conv2d():
    torch = pytest.importorskip("torch")
    F = torch.nn.functional
    ws = oracle.init_weights(0)
    x = oracle.synth_lowlight(40, 56, 6)
    y, A = oracle.zero_dce(x, ws, return_maps=True)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).double()
    xt = t(x.transpose(2, 0, 1))[None]
    cv = lambda z, i: F.conv2d(z, t(ws[i][0]), t(ws[i][1]), padding=1)
    x1 = F.relu(cv(xt, 0)); x2 = F.relu(cv(x1, 1)); x3 = F.relu(cv(x2, 2)); x4 = F.relu(cv(x3, 3))
    x5 = F.relu(cv(torch.cat([x3, x4], 1), 4)); x6 = F.relu(cv(torch.cat([x2, x5], 1), 5))
    At = torch.tanh(cv(torch.cat([x1, x6], 1), 6))
    yt = xt.clone()
    for n in range(8):
        yt = yt + At[:, 3 * n:3 * n + 3] * yt * (1 - yt)
    assert np.abs(A - At[0].permute(1, 2, 0).numpy()).max() < 1e-6
    assert np.abs(y - yt[0].permute(1, 2, 0).numpy()).max() < 1e-6


def test_dce_fp32_mode_is_close_to_fp64():
    ws = oracle.init_weights(1)
    x = oracle.synth_lowlight(32, 40, 7)
    assert np.abs(oracle.zero_dce(x, ws, dtype=np.float32) - oracle.zero_dce(x, ws)).max() < 5e-6


def test_dce_banded_evaluation_is_exact():
    ws = oracle.init_weights(2)
    x = oracle.synth_lowlight(50, 30, 8)
    assert np.array_equal(oracle.zero_dce(x, ws, band=16), oracle.zero_dce(x, ws))
