"""Stress matrix for the fused Zero-DCE chains (run on a B200): for many batch sizes / frame sizes the fused result
must equal the one-launch-per-layer result bitwise, each case in its own process (a trap kills the context).
    python tools/stress_dce.py
"""
import sys, subprocess
CASE = r'''
import sys, torch
sys.path.insert(0, ".")
import low_light_image_enhancement_b200 as L
from low_light_image_enhancement_b200.api import _enhancer
from low_light_image_enhancement_b200.synth import synth_lowlight_batch
n, h, w, reps = map(int, sys.argv[1:5])
x = synth_lowlight_batch(min(n, 4), h, w, seed=1, device=torch.device("cuda", 0))
x = x.repeat((n + 3) // 4, 1, 1, 1)[:n].contiguous()
e = _enhancer(0); wts = L.init_dce_weights(0)
e.debug_dce_mode(keep_all=False, unfused=True)
ref = L.enhance(x, "zero_dce", dict(weights=wts, precision="bf16")).clone()
e.debug_dce_mode(keep_all=False, unfused=False)
ok = True
for i in range(reps):
    y = L.enhance(x, "zero_dce", dict(weights=wts, precision="bf16"))
    ok = ok and bool(torch.equal(y, ref))
torch.cuda.synchronize()
print("ok" if ok else "MISMATCH")
'''
cases = [(n, 1080, 1920, 3) for n in list(range(1, 25))] + [(64, 1080, 1920, 5), (32, 1080, 1920, 5), (5, 720, 1280, 3), (7, 361, 1000, 3), (3, 2160, 3840, 3), (9, 607, 333, 3)]
bad = 0
for c in cases:
    r = subprocess.run([sys.executable, "-c", CASE] + [str(v) for v in c], capture_output=True, text=True, timeout=300)
    res = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else "FAIL rc=%d" % r.returncode
    if res != "ok": bad += 1
    print(c, res, flush=True)
print("BAD", bad)
