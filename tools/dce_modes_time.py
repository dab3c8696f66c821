"""Per-kernel CUDA-event times of the bf16 Zero-DCE path, fused chains vs one launch per layer."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import low_light_image_enhancement_b200 as L
from low_light_image_enhancement_b200.api import _enhancer
from low_light_image_enhancement_b200.synth import synth_lowlight_batch
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 16
enh = _enhancer(0)
x = synth_lowlight_batch(frames, 1080, 1920, seed=1, device="cuda"); y = torch.empty_like(x)
p = dict(weights=L.init_dce_weights(0), precision="bf16")
for unf in (True, False):
    enh.debug_dce_mode(unfused=unf)
    for _ in range(2): enh.enhance(x, "zero_dce", p, out=y)
    torch.cuda.synchronize(); enh.profile_begin()
    for _ in range(3): enh.enhance(x, "zero_dce", p, out=y)
    prof = enh.profile_end()
    tot = sum(v[1] for v in prof.values()) / 3
    print("unfused" if unf else "fused", f"{frames} frames: {tot:.3f} ms/step", {k: round(v[1] / 3, 3) for k, v in prof.items()}, flush=True)
