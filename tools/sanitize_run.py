"""One small invocation of every kernel family, for compute-sanitizer (memcheck / racecheck / synccheck /
initcheck on small shapes).  Every result is still compared with the oracle so that a sanitizer-clean run is also a
correct one.

    compute-sanitizer --tool memcheck python tools/sanitize_run.py [families...]
families: msrcr direct finish tiled dce_bf16 dce_fp32 dce_simt u8 sharded   (default: all)
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import oracle  # noqa: E402
import low_light_image_enhancement_b200 as L  # noqa: E402


def main():
    fams = sys.argv[1:] or ["msrcr", "direct", "finish", "tiled", "dce_bf16", "dce_fp32", "dce_simt", "u8", "sharded"]
    h, w = 72, 136
    x = oracle.synth_lowlight(h, w, 5)
    xd = torch.from_numpy(x).cuda()
    blob = oracle.pack_weights(oracle.init_weights(0))
    ws = oracle.init_weights(0)

    def close(name, y, ref, tol):
        err = float(np.abs(np.asarray(y, np.float64) - ref).max())
        print(f"{name:28s} max-abs {err:.2e} (tol {tol:g})", flush=True)
        assert err <= tol, name

    if "msrcr" in fams:
        close("msrcr pyramid", L.enhance(xd, "msrcr").cpu().numpy(), oracle.msrcr(x), 1e-4)
        xc = torch.from_numpy(np.ascontiguousarray(x.transpose(2, 0, 1))[None]).cuda()
        close("msrcr pyramid NCHW", L.enhance(xc, "msrcr", dict(layout="NCHW"))[0].permute(1, 2, 0).cpu().numpy(), oracle.msrcr(x), 1e-4)
        p5 = dict(sigmas=(13.0, 20.0, 60.0, 150.0), weights=(0.25,) * 4)
        close("msrcr pyramid 4 scales", L.enhance(xd, "msrcr", p5).cpu().numpy(), oracle.msrcr(x, p5), 1e-4)
    if "direct" in fams:
        close("msrcr direct", L.enhance(xd, "msrcr", dict(blur="direct")).cpu().numpy(), oracle.msrcr(x), 1e-4)
    if "finish" in fams:
        for f in ("minmax", "meanstd", "percentile"):
            prm = dict(finish=f, finish_k=1.5)
            close(f"msrcr finish {f}", L.enhance(xd, "msrcr", prm).cpu().numpy(), oracle.msrcr(x, prm), 5e-3)
    if "tiled" in fams:
        xt = oracle.synth_lowlight(200, 136, 6)
        parts = L.tile_rows(200, 2)
        shards = [torch.from_numpy(xt[r0:r0 + r]).cuda().contiguous() for r0, r in parts]
        y = np.concatenate([o.cpu().numpy() for o in L.msrcr_tiled(shards)], 0)
        close("msrcr tiled (2 shards)", y, oracle.msrcr(xt), 1e-4)
        y = np.concatenate([o.cpu().numpy() for o in L.msrcr_tiled(shards, dict(finish="percentile", finish_k=1.5))], 0)
        close("msrcr tiled percentile", y, oracle.msrcr(xt, dict(finish="percentile", finish_k=1.5)), 5e-3)
    ref = oracle.zero_dce(x, ws)
    if "dce_bf16" in fams:
        y = L.enhance(xd, "zero_dce", dict(weights=blob, precision="bf16")).cpu().numpy()
        close("zero_dce bf16 (fused chains)", y, ref, 2e-2)
        xs = torch.from_numpy(np.stack([x, x[::-1].copy()])).cuda()
        yb = L.enhance(xs, "zero_dce", dict(weights=blob, precision="bf16"))
        assert torch.equal(yb[0].cpu(), torch.from_numpy(y))
    if "dce_fp32" in fams:
        close("zero_dce fp32 (split bf16x3)", L.enhance(xd, "zero_dce", dict(weights=blob, precision="fp32")).cpu().numpy(), ref, 1e-4)
    if "dce_simt" in fams:
        close("zero_dce fp32_simt", L.enhance(xd, "zero_dce", dict(weights=blob, precision="fp32_simt")).cpu().numpy(), ref, 1e-4)
    if "u8" in fams:
        xb = oracle.to_u8(x)
        y = L.enhance(torch.from_numpy(xb).cuda(), "msrcr").cpu().numpy()
        assert np.abs(y.astype(int) - oracle.msrcr_u8(xb)[0].astype(int)).max() <= 1
        y = L.enhance(torch.from_numpy(xb).cuda(), "zero_dce", dict(weights=blob, precision="bf16")).cpu().numpy()
        assert np.abs(y.astype(int) - oracle.zero_dce_u8(xb, ws)[0].astype(int)).max() <= 3
        print("u8 paths ok", flush=True)
    if "sharded" in fams:
        xs = np.stack([x, x[::-1].copy(), x[:, ::-1].copy()])
        y = L.enhance(xs, "msrcr", devices=[0, 0])
        close("sharded host msrcr", y[0], oracle.msrcr(x), 1e-4)
        outs = L.enhance([torch.from_numpy(xs[:2]).cuda(), torch.from_numpy(xs[2:]).cuda()], "zero_dce",
                         dict(weights=blob, precision="bf16"), devices=[0, 0])
        close("sharded resident zero_dce", outs[0][0].cpu().numpy(), ref, 2e-2)
    torch.cuda.synchronize()
    print("sanitize_run: all families done", flush=True)


if __name__ == "__main__":
    main()
