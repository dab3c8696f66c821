"""Developer check on a B200: parity numbers and rough timings for every CUDA path (prints, no asserts)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oracle
import low_light_image_enhancement_b200 as L

def timeit(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

def main():
    which = sys.argv[1:] or ["msrcr", "dce"]
    print(torch.cuda.get_device_name(0), torch.cuda.get_device_properties(0).multi_processor_count, "SMs")
    if "msrcr" in which:
        x = oracle.synth_lowlight(400, 600, 0)
        ref = oracle.msrcr(x, mode="fft")
        for blur in ("direct", "pyramid"):
            y = L.enhance(x, "msrcr", dict(blur=blur))
            print(f"msrcr {blur} 600x400 NHWC max-abs {np.abs(y - ref).max():.3e} mean {np.abs(y - ref).mean():.3e}")
            xc = np.ascontiguousarray(x.transpose(2, 0, 1))[None]
            yc = L.enhance(xc, "msrcr", dict(blur=blur, layout="NCHW"))
            print(f"msrcr {blur} 600x400 NCHW max-abs {np.abs(yc[0].transpose(1, 2, 0) - ref).max():.3e}")
        for (h, w, seed) in ((1080, 1920, 5), (2160, 3840, 7)):
            x = oracle.synth_lowlight(h, w, seed)
            t = time.time(); ref = oracle.msrcr(x, mode="fft"); tcpu = time.time() - t
            xd = torch.from_numpy(x).cuda()[None]
            yd = L.enhance(xd, "msrcr")
            d = np.abs(yd[0].cpu().numpy() - ref)
            ms = timeit(lambda: L.enhance(xd, "msrcr"))
            print(f"msrcr pyramid {w}x{h} max-abs {d.max():.3e} mean {d.mean():.3e}  gpu {ms:.3f} ms "
                  f"({h*w/ms/1e3:.1f} Mpx/s)  oracle fft {tcpu:.2f} s")
        xb = torch.from_numpy(np.stack([oracle.synth_lowlight(2160, 3840, 100 + i) for i in range(4)])).cuda()
        xb = xb.repeat(8, 1, 1, 1)
        ms = timeit(lambda: L.enhance(xb, "msrcr"), 3)
        print(f"msrcr pyramid 32x4K: {ms:.3f} ms  {32*2160*3840/ms/1e3:.1f} Mpx/s  -> {108*32*2160*3840/ms/1e6:.1f} GB/s algorithmic(108B/px)")
        xd = xb[:1].contiguous()
        ms = timeit(lambda: L.enhance(xd, "msrcr", dict(blur="direct")), 2)
        print(f"msrcr direct 1x4K: {ms:.3f} ms  {2160*3840/ms/1e3:.1f} Mpx/s")
    if "dce" in which:
        ws = oracle.init_weights(0); blob = oracle.pack_weights(ws)
        for (h, w, seed) in ((135, 240, 1), (270, 481, 2)):
            x = oracle.synth_lowlight(h, w, seed)
            ref, A = oracle.zero_dce(x, ws, return_maps=True)
            for prec in ("fp32", "bf16"):
                try:
                    y, m = L.enhance(x, "zero_dce", dict(weights=blob, precision=prec, return_maps=True))
                except L.LlieError as e:
                    print("dce", prec, "failed:", e); continue
                mse = float(np.mean((y.astype(np.float64) - ref) ** 2))
                print(f"dce {prec} {w}x{h} out max-abs {np.abs(y - ref).max():.3e} maps max-abs {np.abs(m - A).max():.3e} "
                      f"PSNR {10*np.log10(1/max(mse,1e-30)):.1f} dB")
        if "tcdebug" in which:
            sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
            from bf16_emulation import dce_activations_bf16
            from low_light_image_enhancement_b200.api import _enhancer
            for (h, w, seed) in ((20, 40, 4), (70, 200, 5)):
                x = oracle.synth_lowlight(h, w, seed) * 6
                acts, Aem = dce_activations_bf16(x, ws)
                e = _enhancer(0)
                e.debug_dce_mode(keep_all=True, unfused="unfused" in which)
                y, m = L.enhance(torch.from_numpy(x).cuda(), "zero_dce", dict(weights=blob, precision="bf16", return_maps=True))
                for i in range(6):
                    g = e.debug_dce_activation(i + 1, 0, h, w)
                    d = np.abs(g - acts[i]); k = np.unravel_index(d.argmax(), d.shape)
                    print(f"  tc {w}x{h} x{i+1}: max-abs {d.max():.3e} (ref max {np.abs(acts[i]).max():.3f}) at c,y,x={k} frac>1e-2 {(d>1e-2).mean():.4f}")
                mm = m.cpu().numpy(); d = np.abs(mm - Aem.transpose(1, 2, 0))
                print(f"  tc {w}x{h} maps: max-abs {d.max():.3e}")
        x = torch.from_numpy(oracle.synth_lowlight(1080, 1920, 3)).cuda()[None]
        for prec, nb in (("fp32", 1), ("bf16", 1), ("bf16", 8)):
            xb = x.repeat(nb, 1, 1, 1)
            try:
                ms = timeit(lambda: L.enhance(xb, "zero_dce", dict(weights=blob, precision=prec)), 3)
            except L.LlieError as e:
                print("dce", prec, "failed:", e); continue
            px = nb * 1080 * 1920
            print(f"dce {prec} {nb}x1080p: {ms:.3f} ms  {px/ms/1e3:.1f} Mpx/s  {156672*px/ms/1e9:.1f} TFLOP/s(tensor-alg)")

if __name__ == "__main__":
    main()
