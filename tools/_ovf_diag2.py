import sys, os, numpy as np, torch
sys.path.insert(0, ".")
import low_light_image_enhancement_b200 as L
from low_light_image_enhancement_b200.api import _enhancer
from low_light_image_enhancement_b200.synth import synth_lowlight_batch
n = int(sys.argv[1])
h, w = 1080, 1920
x = synth_lowlight_batch(min(n, 4), h, w, seed=1, device=torch.device("cuda", 0))
x = x.repeat((n + 3) // 4, 1, 1, 1)[:n].contiguous()
e = _enhancer(0); wts = L.init_dce_weights(0)
prm = dict(weights=wts, precision="bf16")
KS = (1, 2, 3, 5)
def acts(unfused):
    e.debug_dce_mode(keep_all=False, unfused=unfused)
    y = L.enhance(x, "zero_dce", prm)
    torch.cuda.synchronize()
    return [np.stack([e.debug_dce_activation(k, f, h, w) for f in range(n)]) for k in KS] + [y.cpu().numpy()]
ref = acts(True)
for rep in range(4):
    got = acts(False)
    for i, k in enumerate(list(KS) + [7]):
        bad = got[i] != ref[i]
        if bad.any():
            if k < 7:
                f, c, yy, xx = np.nonzero(bad)
            else:
                f, yy, xx, c = np.nonzero(bad)
            d = np.abs(got[i].astype(np.float64) - ref[i])
            print(f"rep={rep} first bad tensor x{k}: {bad.sum()} values, max diff {d.max():.4g}; frames {np.unique(f)}; rows {yy.min()}..{yy.max()} n_rows {len(np.unique(yy))} "
                  f"first rows {np.unique(yy)[:24]}; cols {xx.min()}..{xx.max()} strips(126) {np.unique(xx // 126)}; ch {np.unique(c)[:12]}; rows%6 {np.bincount(yy % 6, minlength=6)}")
            for ff in np.unique(f)[:3]:
                m = f == ff
                print("   frame", ff, "rows", np.unique(yy[m])[:30], "cols", xx[m].min(), xx[m].max())
            break
    else:
        print(f"rep={rep}: fused == unfused")
