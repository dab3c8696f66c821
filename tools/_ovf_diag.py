import sys, os, numpy as np, torch
sys.path.insert(0, ".")
import low_light_image_enhancement_b200 as L
from low_light_image_enhancement_b200.api import _enhancer
from low_light_image_enhancement_b200.synth import synth_lowlight_batch
mode, n = sys.argv[1], int(sys.argv[2])
h, w = 1080, 1920
x = synth_lowlight_batch(min(n, 4), h, w, seed=1, device=torch.device("cuda", 0))
x = x.repeat((n + 3) // 4, 1, 1, 1)[:n].contiguous()
e = _enhancer(0); wts = L.init_dce_weights(0)
prm = dict(weights=wts, precision="bf16")
def acts(unfused):
    e.debug_dce_mode(keep_all=True, unfused=unfused)
    y = L.enhance(x, "zero_dce", prm)
    torch.cuda.synchronize()
    return [np.stack([e.debug_dce_activation(k, f, h, w) for f in range(n)]) for k in range(1, 7)] + [y.cpu().numpy()]
if mode == "ref":
    np.savez("/tmp/ovf_ref.npz", *acts(True))
    print("ref saved")
else:
    ref = np.load("/tmp/ovf_ref.npz")
    ref = [ref[f"arr_{i}"] for i in range(7)]
    for unf in (True, False):
        for rep in range(3):
            got = acts(unf)
            for k in range(7):
                bad = got[k] != ref[k]
                if bad.any():
                    if k < 6:
                        f, c, yy, xx = np.nonzero(bad)
                    else:
                        f, yy, xx, c = np.nonzero(bad)
                    print(f"unfused={unf} rep={rep} first bad tensor x{k+1}: {bad.sum()} values; frames {np.unique(f)}; rows {yy.min()}..{yy.max()} "
                          f"unique rows {np.unique(yy)[:40]}; cols {xx.min()}..{xx.max()} strips(126) {np.unique(xx // 126)[:20]} strips(124) {np.unique(xx//124)[:20]}; ch {np.unique(c)[:8]}")
                    d = np.abs(got[k].astype(np.float64) - ref[k])
                    print("   max abs diff", d.max(), "rows%6:", np.bincount(yy % 6, minlength=6))
                    break
            else:
                print(f"unfused={unf} rep={rep}: all equal")
