"""Device-side synthetic low-light frames for benchmarks (SURVEY.md 8(d) 'synthetic input').

Same recipe as the test oracle's NumPy generator -- smooth sinusoidal base, hard-edged blended
rectangles, x = 0.12 * base^2.2, signal-dependent + read noise, clip -- but evaluated with torch ops
on the GPU so that a 32-frame 4K batch is generated in milliseconds.  It is measurement plumbing,
not part of the enhancement path; parity tests always feed the CUDA path and the oracle the same
bytes (the oracle receives a D2H copy of these frames when the two must agree)."""
from __future__ import annotations

import math

import torch


def synth_lowlight_batch(n: int, h: int, w: int, seed: int = 0, device="cuda", layout: str = "NHWC") -> torch.Tensor:
    """float32 (n, h, w, 3) (or (n, 3, h, w)) in [0, 1]; mean about 0.025 (genuinely dark)."""
    dev = torch.device(device)
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    out = torch.empty((n, h, w, 3), dtype=torch.float32, device=dev)
    yy = torch.linspace(0.0, 1.0, h, device=dev)[:, None]
    xx = torch.linspace(0.0, 1.0, w, device=dev)[None, :]
    gd = torch.Generator(device=dev)
    for i in range(n):
        base = torch.empty((h, w, 3), dtype=torch.float32, device=dev)
        for c in range(3):
            acc = torch.zeros((h, w), dtype=torch.float32, device=dev)
            for _ in range(6):
                fy, fx = (torch.rand(2, generator=g) * 7.5 + 0.5).tolist()
                amp = float(torch.rand(1, generator=g) * 0.7 + 0.3)
                py, px = (torch.rand(2, generator=g) * 2 * math.pi).tolist()
                acc += amp * torch.sin(2 * math.pi * fy * yy + py) * torch.sin(2 * math.pi * fx * xx + px)
            acc -= acc.min()
            acc /= acc.max().clamp_min(1e-12)
            base[:, :, c] = acc
        for _ in range(20):
            y0 = int(torch.randint(0, h, (1,), generator=g))
            x0 = int(torch.randint(0, w, (1,), generator=g))
            hh = int(torch.randint(max(2, h // 16), max(3, h // 3), (1,), generator=g))
            ww = int(torch.randint(max(2, w // 16), max(3, w // 3), (1,), generator=g))
            col = torch.rand(3, generator=g).to(dev)
            base[y0:y0 + hh, x0:x0 + ww, :] = 0.5 * base[y0:y0 + hh, x0:x0 + ww, :] + 0.5 * col
        x = 0.12 * base.pow(2.2)
        gd.manual_seed(int(seed) * 1000003 + i)
        n1 = torch.randn(x.shape, generator=gd, device=dev)
        n2 = torch.randn(x.shape, generator=gd, device=dev)
        out[i] = (x + 0.01 * n1 * x.sqrt() + 0.002 * n2).clamp_(0.0, 1.0)
    if layout == "NCHW":
        out = out.permute(0, 3, 1, 2).contiguous()
    return out
