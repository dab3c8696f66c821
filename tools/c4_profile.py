"""Per-device kernel times of the row-tiled 16384^2 MSRCR (c4): where the time between tile_t0 and tile_t1 goes.
    python tools/c4_profile.py        (uses every visible GPU)
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import low_light_image_enhancement_b200 as L
from low_light_image_enhancement_b200.api import msrcr_tiled_last_ms, _tile_enhancer
from low_light_image_enhancement_b200.synth import synth_lowlight_batch
H = int(os.environ.get("C4_H", "16384"))
W = int(os.environ.get("C4_W", "16384"))   # C4_H=4096 on 2 GPUs gives the 2048-row shards of the 8-GPU run
g = torch.cuda.device_count()
parts = L.tile_rows(H, g)
shards, outs = [], []
for d, (r0, r) in enumerate(parts):
    bands = [synth_lowlight_batch(1, min(2048, r - b0), W, seed=4000 + 97 * d + b0, device=torch.device("cuda", d))[0]
             for b0 in range(0, r, 2048)]
    shards.append(torch.cat(bands, 0).contiguous()); outs.append(torch.empty_like(shards[-1]))
for _ in range(3):
    L.msrcr_tiled(shards, outs=outs)
enh = [_tile_enhancer(d, 0) for d in range(g)]
ms = []
for _ in range(5):
    L.msrcr_tiled(shards, outs=outs); ms.append(msrcr_tiled_last_ms(shards))
print("tile ms per device (5 calls):", [[round(v, 3) for v in m] for m in ms])
for e in enh: e.profile_begin()
L.msrcr_tiled(shards, outs=outs)
print("profiled call tile ms:", [round(v, 3) for v in msrcr_tiled_last_ms(shards)])
for d, e in enumerate(enh):
    p = e.profile_end()
    tot = sum(v[1] for v in p.values())
    print(f"dev {d}: kernels {tot:.3f} ms in {sum(v[0] for v in p.values())} launches", {k: round(v[1], 3) for k, v in sorted(p.items())})
