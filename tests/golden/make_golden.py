"""Regenerates the committed golden vectors from the oracle (python tests/golden/make_golden.py).

The reference ships no fixtures (/root/reference/README.md:1-2), so these vectors pin the
self-written oracle against accidental change and give the GPU tests fixed targets.
Inputs are stored too, so the vectors do not depend on NumPy's RNG stream staying stable.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle  # noqa: E402


def main():
    x = oracle.synth_lowlight(96, 128, seed=11)
    y = oracle.msrcr(x, mode="direct")
    np.savez_compressed(os.path.join(HERE, "msrcr_96x128.npz"), x=x, y=y)
    # a second MSRCR case with non-default scales/gain (sigma = 6 is below the cascade's range: direct blur only)
    p = dict(sigmas=(6.0, 33.0, 120.0), weights=(0.5, 0.3, 0.2), gain=0.8, offset=120.0)
    x2 = oracle.synth_lowlight(80, 112, seed=12)
    np.savez_compressed(os.path.join(HERE, "msrcr_80x112_custom.npz"), x=x2, y=oracle.msrcr(x2, p, mode="direct"),
                        sigmas=np.array(p["sigmas"]), weights=np.array(p["weights"]), gain=p["gain"], offset=p["offset"])
    # a third one whose scales all sit inside the decimated cascade's range (sigma >= 12), on both levels
    p3 = dict(sigmas=(13.0, 33.0, 120.0), weights=(0.5, 0.3, 0.2), gain=0.8, offset=120.0)
    x3 = oracle.synth_lowlight(88, 120, seed=14)
    np.savez_compressed(os.path.join(HERE, "msrcr_88x120_custom.npz"), x=x3, y=oracle.msrcr(x3, p3, mode="direct"),
                        sigmas=np.array(p3["sigmas"]), weights=np.array(p3["weights"]), gain=p3["gain"], offset=p3["offset"])
    ws = oracle.init_weights(seed=0)
    xd = oracle.synth_lowlight(48, 64, seed=13)
    yd, A = oracle.zero_dce(xd, ws, return_maps=True)
    np.savez_compressed(os.path.join(HERE, "zero_dce_48x64.npz"), x=xd, y=yd, maps=A.astype(np.float32),
                        weights=oracle.pack_weights(ws))
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
