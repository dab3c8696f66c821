"""Short, deterministic launch sequence for ncu: W warm-up calls + 1 profiled call of one workload.

    python tools/prof_run.py dce [frames]     Zero-DCE bf16 on `frames` 1080p frames (default 8)
    python tools/prof_run.py msrcr [frames]   MSRCR on `frames` 4K frames (default 8); msrcr_u8: the same on uint8 images
    python tools/prof_run.py dce32            Zero-DCE fp32 on one 1080p frame
Prints the number of kernel launches per call so that `ncu -s` can skip the warm-up.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import low_light_image_enhancement_b200 as L  # noqa: E402
from low_light_image_enhancement_b200.api import _enhancer  # noqa: E402
from low_light_image_enhancement_b200.synth import synth_lowlight_batch  # noqa: E402


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "dce"
    frames = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    warm = int(os.environ.get("PROF_WARMUP", "1"))
    dev = torch.device("cuda", 0)
    enh = _enhancer(0)
    if what in ("msrcr", "msrcr_u8"):
        x = synth_lowlight_batch(frames, 2160, 3840, seed=3000, device=dev)
        if what == "msrcr_u8":
            x = (x * 255.0).round().clamp(0, 255).to(torch.uint8)
        method, params = "msrcr", None
    elif what == "dce32":
        x = synth_lowlight_batch(1, 1080, 1920, seed=1000, device=dev)
        method, params = "zero_dce", dict(weights=L.init_dce_weights(0), precision="fp32")
    else:
        x = synth_lowlight_batch(frames, 1080, 1920, seed=2000, device=dev)
        method, params = "zero_dce", dict(weights=L.init_dce_weights(0), precision="bf16")
    y = torch.empty_like(x)
    l0 = enh.launch_count()
    for _ in range(warm):
        enh.enhance(x, method, params, out=y)
    torch.cuda.synchronize()
    l1 = enh.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    enh.enhance(x, method, params, out=y)
    e1.record()
    torch.cuda.synchronize()
    per_call = enh.launch_count() - l1
    print(f"{what}: {frames} frame(s), {per_call} launches per call, warm-up launches {l1 - l0}, "
          f"last call {e0.elapsed_time(e1):.3f} ms", flush=True)


if __name__ == "__main__":
    main()
