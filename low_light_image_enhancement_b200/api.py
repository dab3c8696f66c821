"""The public operator: enhance(images, method, params) (BASELINE.json#north_star).

The reference repository has no operator or plugin interface to mirror
(/root/reference/README.md:1-2 is its entire content), so the interface is the one
north_star specifies: one function accepting NumPy arrays or PyTorch tensors.
"""
from __future__ import annotations

import ctypes as C
import math
import threading
import zlib
from typing import Any, Sequence

import numpy as np
import torch

from . import _lib
from ._lib import LlieError, MsrcrParams, DceParams

DCE_LAYERS = ((32, 3), (32, 32), (32, 32), (32, 32), (32, 64), (32, 64), (24, 64))
DCE_NUM_PARAMS = sum(co * ci * 9 + co for co, ci in DCE_LAYERS)


_MSRCR_DEFAULTS: dict | None = None


def default_msrcr_params() -> dict:
    """SURVEY.md 9.1 defaults (filled by the library so both sides agree; fetched once)."""
    global _MSRCR_DEFAULTS
    if _MSRCR_DEFAULTS is None:
        _MSRCR_DEFAULTS = _fetch_msrcr_defaults()
    return dict(_MSRCR_DEFAULTS)


def _fetch_msrcr_defaults() -> dict:
    p = MsrcrParams()
    p.struct_size = C.sizeof(MsrcrParams)   # the caller states its struct's size; the library writes no further
    _lib.check(_lib.load().llie_msrcr_default_params(C.byref(p)))
    n = p.n_scales
    return dict(sigmas=tuple(p.sigma[:n]), weights=tuple(p.weight[:n]), truncate=p.truncate, alpha=p.alpha,
                beta=p.beta, gain=p.gain, offset=p.offset, in_scale=p.in_scale, in_bias=p.in_bias, blur="pyramid",
                finish="none", finish_k=p.finish_k)


def default_dce_params() -> dict:
    return dict(iters=8, curve_sign=1, precision="fp32", return_maps=False, weights=None)


def init_dce_weights(seed: int = 0, last_gain: float = 1.0) -> np.ndarray:
    """Random-init DCE-Net weights as the packed 79 416-float blob (SURVEY.md 8(d) 'DCE weights':
    He-normal W ~ N(0, 2/(9 Cin)), b ~ U(-0.05, 0.05), drawn W1,b1,...,W7,b7 from default_rng(seed))."""
    rng = np.random.default_rng(seed)
    parts = []
    for li, (co, ci) in enumerate(DCE_LAYERS):
        w = rng.normal(0.0, np.sqrt(2.0 / (9.0 * ci)), size=(co, ci, 3, 3)).astype(np.float32)
        b = rng.uniform(-0.05, 0.05, size=(co,)).astype(np.float32)
        if li == len(DCE_LAYERS) - 1 and last_gain != 1.0:
            w = (w * np.float32(last_gain)).astype(np.float32)
        parts += [w.ravel(), b]
    return np.concatenate(parts).astype(np.float32)


_UPSTREAM_KEYS = tuple(f"e_conv{i}" for i in range(1, 8))


def load_dce_state_dict(src, *, strict: bool = True) -> np.ndarray:
    """Packed weight blob from a Zero-DCE checkpoint.

    src : a path to a torch checkpoint (.pth, loaded with weights_only=True on the CPU) or a mapping
          name -> tensor/array.  Expected entries are the public Zero-DCE layer names e_conv1 .. e_conv7
          (".weight" (Cout, Cin, 3, 3) and ".bias" (Cout,)), optionally behind a "module." prefix or inside a
          {"state_dict": ...} wrapper.  The public model applies x + r (x^2 - x), i.e. pass
          params["curve_sign"] = -1 with these weights (north_star's convention is +1).
    Returns the 79 416-float blob W1 b1 ... W7 b7 that enhance(..., params={"weights": blob}) takes.
    """
    if isinstance(src, (str, bytes)) or hasattr(src, "__fspath__"):
        sd = torch.load(src, map_location="cpu", weights_only=True)
    else:
        sd = src
    if isinstance(sd, dict) and "state_dict" in sd and not any(str(k).startswith("e_conv") for k in sd):
        sd = sd["state_dict"]
    flat = {}
    for k, v in sd.items():
        k = str(k)
        while k.startswith("module."):
            k = k[len("module."):]
        flat[k] = v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)
    parts = []
    for name, (co, ci) in zip(_UPSTREAM_KEYS, DCE_LAYERS):
        try:
            w, b = flat.pop(name + ".weight"), flat.pop(name + ".bias")
        except KeyError as e:
            raise ValueError(f"checkpoint has no entry {e.args[0]!r}") from None
        if w.shape != (co, ci, 3, 3) or b.shape != (co,):
            raise ValueError(f"{name}: expected weight {(co, ci, 3, 3)} and bias {(co,)}, got {w.shape} and {b.shape}")
        parts += [np.asarray(w, np.float32).ravel(), np.asarray(b, np.float32).ravel()]
    if strict and flat:
        raise ValueError(f"unexpected entries in the checkpoint: {sorted(flat)[:5]}")
    blob = np.concatenate(parts).astype(np.float32)
    assert blob.size == DCE_NUM_PARAMS
    return blob


def plan_query(sigma: float, truncate: float = 4.0) -> dict:
    """Host-side planner of the pyramid blur (needs no GPU): stride, low-res radius and taps, and the three
    accuracy figures of include/llie.h (l1_error; edge_error / line_error = relative error of the surround of a
    black pixel beside a white half-plane / a 1-pixel white line, for I = 255 x + 1)."""
    lib = _lib.load()
    st, sl, rl, l1 = C.c_int(), C.c_double(), C.c_int(), C.c_double()
    _lib.check(lib.llie_msrcr_plan_query(float(sigma), float(truncate), C.byref(st), C.byref(sl), C.byref(rl), C.byref(l1)))
    e1, ee, el = C.c_double(), C.c_double(), C.c_double()
    _lib.check(lib.llie_msrcr_plan_errors(float(sigma), float(truncate), C.byref(e1), C.byref(ee), C.byref(el)))
    cnt = C.c_int()
    _lib.check(lib.llie_msrcr_plan_taps(float(sigma), float(truncate), None, 0, C.byref(cnt)))
    taps = np.empty(cnt.value, np.float32)
    _lib.check(lib.llie_msrcr_plan_taps(float(sigma), float(truncate), taps.ctypes.data_as(C.POINTER(C.c_float)),
                                        cnt.value, C.byref(cnt)))
    return dict(stride=st.value, sigma_low=sl.value, radius_low=rl.value, l1_error=l1.value, edge_error=ee.value,
                line_error=el.value, taps=taps)


def shard_slice(n: int, rank: int, world: int) -> slice:
    """Contiguous split of n independent images over `world` ranks: rank r gets
    [r*n/world, (r+1)*n/world) with the remainder spread over the first ranks.  No collective."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(n, world)
    start = rank * base + min(rank, rem)
    return slice(start, start + base + (1 if rank < rem else 0))


_IMAGE_DTYPES = (torch.float32, torch.uint8, torch.uint16, torch.bfloat16)   # bfloat16: zero_dce, precision bf16


def _msrcr_struct(params: dict | None) -> MsrcrParams:
    d = default_msrcr_params()
    if params:
        unknown = set(params) - set(d) - {"layout"}
        if unknown:
            raise ValueError(f"unknown msrcr params: {sorted(unknown)}")
        d.update({k: v for k, v in params.items() if k != "layout"})
    sig = [float(s) for s in d["sigmas"]]
    wts = [float(w) for w in d["weights"]]
    if len(sig) != len(wts) or not (1 <= len(sig) <= _lib.MAX_SCALES):
        raise ValueError("sigmas and weights must have equal length in [1, 8]")
    p = MsrcrParams()
    p.struct_size = C.sizeof(MsrcrParams)
    p.n_scales = len(sig)
    for i, (s, w) in enumerate(zip(sig, wts)):
        p.sigma[i], p.weight[i] = s, w
    for k in ("truncate", "alpha", "beta", "gain", "offset", "in_scale", "in_bias"):
        setattr(p, k, float(d[k]))
    blur = d["blur"]
    if blur not in ("pyramid", "direct"):
        raise ValueError("blur must be 'pyramid' or 'direct'")
    p.blur = _lib.BLUR_PYRAMID if blur == "pyramid" else _lib.BLUR_DIRECT
    if d["finish"] not in _lib.FINISH:
        raise ValueError("finish must be 'none', 'minmax', 'meanstd' or 'percentile'")
    p.finish, p.finish_k = _lib.FINISH[d["finish"]], float(d["finish_k"])
    return p


def _dce_struct(d: dict) -> DceParams:
    p = DceParams()
    p.struct_size = C.sizeof(DceParams)
    p.iters = int(d["iters"])
    p.curve_sign = int(d["curve_sign"])
    prec = d["precision"]
    codes = {"fp32": _lib.F32, "bf16": _lib.BF16, "fp32_simt": _lib.F32_SIMT}
    if prec not in codes:
        raise ValueError("precision must be 'fp32', 'bf16' or 'fp32_simt'")
    p.precision = codes[prec]
    return p


def _resolve_layout(shape: Sequence[int], layout: str | None):
    """Returns (n, h, w, layout_code, batched) for (H,W,3) / (N,H,W,3) / (N,3,H,W)."""
    if len(shape) == 3:
        if shape[2] != 3:
            raise ValueError(f"3-D input must be (H, W, 3), got {tuple(shape)}")
        if layout not in (None, "NHWC"):
            raise ValueError("a 3-D image is (H, W, 3); layout must be NHWC")
        return 1, shape[0], shape[1], _lib.NHWC, False
    if len(shape) != 4:
        raise ValueError(f"images must be (H,W,3), (N,H,W,3) or (N,3,H,W), got {tuple(shape)}")
    nhwc_ok, nchw_ok = shape[3] == 3, shape[1] == 3
    if layout is None:
        if nhwc_ok and nchw_ok:
            raise ValueError(f"shape {tuple(shape)} is ambiguous; pass params['layout'] = 'NHWC' or 'NCHW'")
        layout = "NHWC" if nhwc_ok else "NCHW" if nchw_ok else None
    if layout == "NHWC" and nhwc_ok:
        return shape[0], shape[1], shape[2], _lib.NHWC, True
    if layout == "NCHW" and nchw_ok:
        return shape[0], shape[2], shape[3], _lib.NCHW, True
    raise ValueError(f"shape {tuple(shape)} does not match layout {layout}")


class Enhancer:
    """One libllie context on one device plus the staging buffers NumPy callers need.

    Not thread-safe: use one Enhancer per host thread.
    """

    def __init__(self, device: int | None = None):
        self._lib = _lib.load()
        n = C.c_int()
        _lib.check(self._lib.llie_device_count(C.byref(n)))  # raises without a GPU: no CPU fallback
        self.device = torch.cuda.current_device() if device is None else int(device)
        h = C.c_void_p()
        _lib.check(self._lib.llie_create(C.byref(h), self.device))
        self._h = h
        self._weights_key = None
        self._stage: dict[str, torch.Tensor] = {}
        self._copy_stream = None

    def close(self):
        if getattr(self, "_h", None):
            self._lib.llie_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- diagnostics -------------------------------------------------------------------------
    def launch_count(self) -> int:
        c = C.c_uint64()
        _lib.check(self._lib.llie_launch_count(self._h, C.byref(c)))
        return c.value

    def workspace_bytes(self) -> int:
        c = C.c_size_t()
        _lib.check(self._lib.llie_workspace_bytes(self._h, C.byref(c)))
        return c.value

    def profile_begin(self) -> None:
        """Start per-kernel CUDA-event timing of everything this Enhancer launches."""
        _lib.check(self._lib.llie_profile_begin(self._h))

    def profile_end(self) -> dict:
        """Stop timing; returns {kernel_name: (launches, total_ms)}."""
        buf = C.create_string_buffer(1 << 16)
        _lib.check(self._lib.llie_profile_end(self._h, buf, len(buf)))
        out = {}
        for rec in buf.value.decode().split(";"):
            if rec:
                name, cnt, ms = rec.split(":")
                out[name] = (int(cnt), float(ms))
        return out

    def debug_dce_activation(self, which: int, frame: int, h: int, w: int) -> np.ndarray:
        """Test hook: inter-layer activation x{which} (1..6) of the last bf16 forward as fp32 (32, h, w)."""
        out = np.empty((32, h, w), np.float32)
        _lib.check(self._lib.llie_debug_dce_activation(self._h, which, frame, out.ctypes.data_as(C.POINTER(C.c_float))))
        return out

    def debug_dce_mode(self, keep_all: bool = False, unfused: bool = False) -> None:
        """Test hook (bf16 path): keep_all also stores x4/x6 to HBM; unfused runs one launch per layer."""
        _lib.check(self._lib.llie_debug_dce_mode(self._h, (1 if keep_all else 0) | (2 if unfused else 0)))

    # -- weights -----------------------------------------------------------------------------
    def set_dce_weights(self, weights) -> None:
        blob = _as_weight_blob(weights)
        key = (blob.size, zlib.crc32(blob.data))   # by content (318 KB: microseconds), not by address
        if key == self._weights_key:
            return
        _lib.check(self._lib.llie_dce_set_weights(self._h, blob.ctypes.data_as(C.POINTER(C.c_float)), blob.size))
        self._weights_key = key

    # -- device-pointer entry points (zero-copy for torch CUDA tensors) -----------------------
    def msrcr_device(self, x: torch.Tensor, out: torch.Tensor, n, h, w, layout, p: MsrcrParams, stream_ptr):
        fn = {torch.uint8: self._lib.llie_msrcr_u8, torch.uint16: self._lib.llie_msrcr_u16}.get(x.dtype, self._lib.llie_msrcr_f32)
        _lib.check(fn(self._h, x.data_ptr(), out.data_ptr(), n, h, w, layout, C.byref(p), stream_ptr))

    def dce_device(self, x, out, maps, n, h, w, layout, p: DceParams, stream_ptr):
        fn = {torch.uint8: self._lib.llie_dce_forward_u8, torch.uint16: self._lib.llie_dce_forward_u16,
              torch.bfloat16: self._lib.llie_dce_forward_bf16io}.get(x.dtype, self._lib.llie_dce_forward)
        _lib.check(fn(self._h, x.data_ptr(), out.data_ptr(), maps.data_ptr() if maps is not None else None,
                      n, h, w, layout, C.byref(p), stream_ptr))

    def dce_curves_device(self, x, maps, out, n, h, w, layout, p: DceParams, stream_ptr):
        _lib.check(self._lib.llie_dce_curves(self._h, x.data_ptr(), maps.data_ptr(), out.data_ptr(), n, h, w,
                                             layout, C.byref(p), stream_ptr))

    # -- the operator ------------------------------------------------------------------------
    def enhance(self, images, method: str, params: dict | None = None, *, out=None):
        params = dict(params or {})
        if method not in ("msrcr", "zero_dce"):
            raise ValueError(f"method must be 'msrcr' or 'zero_dce', got {method!r}")
        is_np = isinstance(images, np.ndarray)
        if not is_np and not isinstance(images, torch.Tensor):
            raise TypeError("images must be a numpy.ndarray or a torch.Tensor")
        t = torch.from_numpy(images) if is_np else images
        if t.dtype not in _IMAGE_DTYPES or (t.dtype == torch.bfloat16 and method != "zero_dce"):
            raise TypeError(f"images must be float32 (values in [0,1]), uint8 (b / 255), uint16 (b / 65535) or, for "
                            f"zero_dce with precision='bf16', bfloat16 tensors; got {t.dtype} for {method}")
        if not t.is_contiguous():
            raise ValueError("images must be contiguous")
        n, h, w, layout, batched = _resolve_layout(t.shape, params.get("layout"))
        mp, dp, want_maps = self._method_structs(method, params)

        if t.is_cuda:
            if t.device.index != self.device:
                raise ValueError(f"tensor on cuda:{t.device.index}, Enhancer on cuda:{self.device}")
            y = torch.empty_like(t) if out is None else out
            _check_out(y, t)
            maps = None
            if want_maps:
                mshape = (n, h, w, 24) if layout == _lib.NHWC else (n, 24, h, w)
                maps = torch.empty(mshape if batched else mshape[1:], dtype=torch.float32, device=t.device)
            sp = _stream_ptr(torch.cuda.current_stream(t.device))
            if mp is not None:
                self.msrcr_device(t, y, n, h, w, layout, mp, sp)
            else:
                self.dce_device(t, y, maps, n, h, w, layout, dp, sp)
            return (y, maps) if want_maps else y

        # host path: stage through pinned memory in frame chunks, overlapping H2D / compute / D2H
        y_t = self._host_pipeline(t, n, h, w, layout, mp, dp, want_maps, out, is_np)
        return y_t

    def _method_structs(self, method: str, params: dict):
        """(MsrcrParams | None, DceParams | None, want_maps) of one enhance() call; uploads the DCE-Net weights."""
        if method == "msrcr":
            return _msrcr_struct(params), None, False
        d = default_dce_params()
        unknown = set(params) - set(d) - {"layout"}
        if unknown:
            raise ValueError(f"unknown zero_dce params: {sorted(unknown)}")
        d.update({k: v for k, v in params.items() if k != "layout"})
        if d["weights"] is None:
            raise ValueError("zero_dce needs params['weights'] (packed blob or list of (W, b))")
        self.set_dce_weights(d["weights"])
        return None, _dce_struct(d), bool(d["return_maps"])

    def stream(self, frames, method: str, params: dict | None = None, *, depth: int = 3, copy: bool = True):
        """Generator behind enhance_stream(): see there.  Parameters are checked here, before the first frame."""
        params = dict(params or {})
        if method not in ("msrcr", "zero_dce"):
            raise ValueError(f"method must be 'msrcr' or 'zero_dce', got {method!r}")
        if depth < 1:
            raise ValueError("depth must be >= 1")
        mp, dp, want_maps = self._method_structs(method, params)
        if want_maps:
            raise ValueError("enhance_stream does not return curve maps")
        def gen():
            dev = torch.device("cuda", self.device)
            h2d, comp, d2h = (torch.cuda.Stream(dev) for _ in range(3))
            slots, shape0 = [], None   # per slot: pinned in / out, device in / out, events, "is numpy"
            inflight = []              # slot indices in submission order

            def take(si):
                sl = slots[si]
                sl["ev_out"].synchronize()
                y = sl["yp"].clone() if copy else sl["yp"]   # the pinned stage is reused `depth` frames later
                return y.numpy() if sl["is_np"] else y

            i = 0
            for f in frames:
                is_np = isinstance(f, np.ndarray)
                t = torch.from_numpy(f) if is_np else f
                if not isinstance(t, torch.Tensor) or t.is_cuda:
                    raise TypeError("enhance_stream takes host frames (numpy.ndarray or CPU torch.Tensor)")
                if t.dtype not in _IMAGE_DTYPES or not t.is_contiguous():
                    raise TypeError("frames must be contiguous float32, uint8 or uint16")
                if shape0 is None:
                    shape0 = (tuple(t.shape), t.dtype)
                    n, h, w, layout, _ = _resolve_layout(t.shape, params.get("layout"))
                elif (tuple(t.shape), t.dtype) != shape0:
                    raise ValueError(f"all frames of a stream must share shape and dtype: {shape0} vs {(tuple(t.shape), t.dtype)}")
                si = i % depth
                if len(slots) <= si:
                    slots.append(dict(xp=torch.empty(t.shape, dtype=t.dtype, pin_memory=True),
                                      yp=torch.empty(t.shape, dtype=t.dtype, pin_memory=True),
                                      xd=torch.empty(t.shape, dtype=t.dtype, device=dev),
                                      yd=torch.empty(t.shape, dtype=t.dtype, device=dev),
                                      ev_in=torch.cuda.Event(), ev_comp=torch.cuda.Event(), ev_out=torch.cuda.Event(),
                                      is_np=is_np))
                else:
                    yield take(inflight.pop(0))            # frame i - depth: its slot is free once it has been handed out
                sl = slots[si]
                sl["is_np"] = is_np
                src = t
                if not t.is_pinned():
                    sl["xp"].copy_(t)                      # host memcpy into the pinned stage (H2D of frame i-depth is done)
                    src = sl["xp"]
                sl["src"] = src                            # a pinned caller frame stays referenced until its slot is reused
                with torch.cuda.stream(h2d):
                    h2d.wait_event(sl["ev_comp"])          # the compute that last read xd has finished
                    sl["xd"].copy_(src, non_blocking=True)
                    sl["ev_in"].record(h2d)
                comp.wait_event(sl["ev_in"])
                comp.wait_event(sl["ev_out"])              # the D2H that last read yd has finished
                if mp is not None:
                    self.msrcr_device(sl["xd"], sl["yd"], n, h, w, layout, mp, _stream_ptr(comp))
                else:
                    self.dce_device(sl["xd"], sl["yd"], None, n, h, w, layout, dp, _stream_ptr(comp))
                sl["ev_comp"].record(comp)
                with torch.cuda.stream(d2h):
                    d2h.wait_event(sl["ev_comp"])
                    sl["yp"].copy_(sl["yd"], non_blocking=True)
                    sl["ev_out"].record(d2h)
                inflight.append(si)
                i += 1
            while inflight:
                yield take(inflight.pop(0))

        return gen()

    def _pinned(self, name: str, nelem: int, dtype=torch.float32) -> torch.Tensor:
        name = f"{name}.{dtype}"
        buf = self._stage.get(name)
        if buf is None or buf.numel() < nelem:
            buf = torch.empty(nelem, dtype=dtype, pin_memory=True)
            self._stage[name] = buf
        return buf[:nelem]

    def _devbuf(self, name: str, nelem: int, dtype=torch.float32) -> torch.Tensor:
        name = f"{name}.{dtype}"
        buf = self._stage.get(name)
        if buf is None or buf.numel() < nelem:
            buf = torch.empty(nelem, dtype=dtype, device=f"cuda:{self.device}")
            self._stage[name] = buf
        return buf[:nelem]

    def _host_pipeline(self, t, n, h, w, layout, mp, dp, want_maps, out, is_np):
        dev = torch.device("cuda", self.device)
        frame = h * w * 3
        x_flat = t.reshape(-1)
        dt = t.dtype
        if out is None:
            y_host = torch.empty(t.shape, dtype=dt)
        else:
            y_host = torch.from_numpy(out) if isinstance(out, np.ndarray) else out
            _check_out(y_host, t)
        y_flat = y_host.reshape(-1)
        maps_host = None
        if want_maps:
            mshape = (n, h, w, 24) if layout == _lib.NHWC else (n, 24, h, w)
            maps_host = torch.empty(mshape if t.dim() == 4 else mshape[1:], dtype=torch.float32)
        in_pinned, out_pinned = x_flat.is_pinned(), y_flat.is_pinned()
        # chunk: about 64 MiB of input per stage, at least one frame
        per = max(1, min(n, (64 << 20) // (frame * t.element_size())))
        n_chunks = math.ceil(n / per)
        if self._copy_stream is None:
            self._copy_stream = [torch.cuda.Stream(dev) for _ in range(2)]
        h2d, d2h = self._copy_stream
        comp = torch.cuda.current_stream(dev)
        xd = [self._devbuf(f"x{i}", per * frame, dt) for i in range(2)]
        yd = [self._devbuf(f"y{i}", per * frame, dt) for i in range(2)]
        md = [self._devbuf(f"m{i}", per * frame * 8) for i in range(2)] if want_maps else [None, None]
        xp = None if in_pinned else [self._pinned(f"xp{i}", per * frame, dt) for i in range(2)]
        yp = None if out_pinned else [self._pinned(f"yp{i}", per * frame, dt) for i in range(2)]
        ev_in = [torch.cuda.Event() for _ in range(2)]
        ev_comp = [torch.cuda.Event() for _ in range(2)]
        ev_out = [torch.cuda.Event() for _ in range(2)]
        pending = []  # (slot, lo, hi) whose D2H has been issued into a pinned stage
        for ci in range(n_chunks):
            s = ci & 1
            lo, hi = ci * per, min(n, (ci + 1) * per)
            cnt = (hi - lo) * frame
            # The host blocks only where it touches a staging buffer itself.  With pinned input and output it never
            # does: every buffer reuse below is ordered by stream-side event waits, so all chunks are enqueued ahead
            # and H2D(ci) overlaps D2H(ci-2) (a host-side wait here serialised the two directions of the link).
            if ci >= 2 and not out_pinned:
                ev_out[s].synchronize()            # slot s: its previous D2H is done
                self._drain(pending, y_flat, yp, frame, upto_slot=s)
            src = x_flat[lo * frame:hi * frame]
            if not in_pinned:
                if ci >= 2:
                    ev_in[s].synchronize()         # the previous H2D out of xp[s] is done
                xp[s][:cnt].copy_(src)             # pageable -> pinned staging (host memcpy)
                src = xp[s][:cnt]
            with torch.cuda.stream(h2d):
                h2d.wait_event(ev_comp[s])         # compute that last read xd[s] has finished
                xd[s][:cnt].copy_(src, non_blocking=True)
                ev_in[s].record(h2d)
            comp.wait_event(ev_in[s])
            comp.wait_event(ev_out[s])             # previous D2H out of yd[s] has finished
            nb = hi - lo
            if mp is not None:
                self.msrcr_device(xd[s], yd[s], nb, h, w, layout, mp, _stream_ptr(comp))
            else:
                self.dce_device(xd[s], yd[s], md[s], nb, h, w, layout, dp, _stream_ptr(comp))
            ev_comp[s].record(comp)
            with torch.cuda.stream(d2h):
                d2h.wait_event(ev_comp[s])
                dst = y_flat[lo * frame:hi * frame] if out_pinned else yp[s][:cnt]
                dst.copy_(yd[s][:cnt], non_blocking=True)
                if want_maps:
                    maps_host.reshape(-1)[lo * frame * 8:hi * frame * 8].copy_(md[s][:cnt * 8])
                ev_out[s].record(d2h)
            if not out_pinned:
                pending.append((s, lo, hi))
        d2h.synchronize()
        self._drain(pending, y_flat, yp, frame, upto_slot=None)
        y = y_host.numpy() if is_np else y_host
        if want_maps:
            return y, (maps_host.numpy() if is_np else maps_host)
        return y

    @staticmethod
    def _drain(pending, y_flat, yp, frame, upto_slot):
        keep = []
        for (s, lo, hi) in pending:
            if upto_slot is None or s == upto_slot:
                y_flat[lo * frame:hi * frame].copy_(yp[s][:(hi - lo) * frame])
            else:
                keep.append((s, lo, hi))
        pending[:] = keep


def _stream_ptr(stream: torch.cuda.Stream) -> int:
    """cudaStream_t for the C ABI.  torch's default stream is the NULL stream, which llie.h reserves
    for "the ctx's own stream, synchronous"; pass CUDA's explicit legacy-stream handle instead so the
    work stays ordered with the caller's torch ops."""
    p = stream.cuda_stream
    return p if p else 0x1  # cudaStreamLegacy


def _check_out(y: torch.Tensor, t: torch.Tensor) -> None:
    if y.shape != t.shape or y.dtype != t.dtype or not y.is_contiguous() or y.device != t.device:
        raise ValueError("out must match images in shape, dtype, device and be contiguous")
    if y.data_ptr() == t.data_ptr():
        raise ValueError("out must not alias images")


def _as_weight_blob(weights) -> np.ndarray:
    if isinstance(weights, torch.Tensor):
        weights = weights.detach().cpu().numpy()
    if isinstance(weights, np.ndarray):
        blob = np.ascontiguousarray(weights, dtype=np.float32).ravel()
    else:  # list of (W, b)
        blob = np.concatenate([np.concatenate([np.asarray(w, np.float32).ravel(), np.asarray(b, np.float32).ravel()])
                               for w, b in weights]).astype(np.float32)
    if blob.size != DCE_NUM_PARAMS:
        raise ValueError(f"DCE-Net weight blob must hold {DCE_NUM_PARAMS} floats, got {blob.size}")
    return blob


_tls = threading.local()


def _enhancer(device: int) -> Enhancer:
    cache = getattr(_tls, "cache", None)
    if cache is None:
        cache = _tls.cache = {}
    e = cache.get(device)
    if e is None:
        e = cache[device] = Enhancer(device)
    return e


def _tile_enhancer(device: int, slot: int) -> Enhancer:
    """Extra contexts for row-tiled MSRCR: one per shard, also when several shards share a device."""
    cache = getattr(_tls, "tile_cache", None)
    if cache is None:
        cache = _tls.tile_cache = {}
    e = cache.get((device, slot))
    if e is None:
        e = cache[(device, slot)] = Enhancer(device)
    return e


def tile_rows(h: int, parts: int) -> list[tuple[int, int]]:
    """Row partition of an h-row image into `parts` shards [(row0, rows), ...] for msrcr_tiled():
    equal shards with interior boundaries on multiples of 4 (the stride of the first cascade level)."""
    if parts <= 0 or h <= 0:
        raise ValueError(f"bad partition {h} rows / {parts} parts")
    bounds = [0] + [min(h, ((h * i // parts) + 3) // 4 * 4) for i in range(1, parts)] + [h]
    out = [(bounds[i], bounds[i + 1] - bounds[i]) for i in range(parts)]
    if any(r <= 0 for _, r in out):
        raise ValueError(f"image of {h} rows is too small for {parts} shards")
    return out


def msrcr_tiled(shards: Sequence[torch.Tensor], params: dict | None = None, *, outs: Sequence[torch.Tensor] | None = None):
    """MSRCR of ONE image held as horizontal shards (BASELINE.json configs[4]).

    shards : CUDA float32 tensors of shape (rows_d, W, 3), top to bottom; shard d lives on the device
             that processes it.  Either every shard is on its own device (halo rows travel as NVLink
             peer-to-peer copies, no NCCL) or all are on one device ("virtual devices", used by the
             single-GPU tests).  Interior boundaries must be multiples of 4 rows (see tile_rows()).
    Returns the enhanced shards (same shapes/devices).  Synchronous: returns when every device is done.
    """
    if len(shards) == 0:
        return []
    for t in shards:
        if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32 and t.dim() == 3
                and t.shape[2] == 3 and t.is_contiguous()):
            raise TypeError("msrcr_tiled shards must be contiguous CUDA float32 tensors of shape (rows, W, 3)")
    w = int(shards[0].shape[1])
    if any(int(t.shape[1]) != w for t in shards):
        raise ValueError("all shards must have the same width")
    outs = [torch.empty_like(t) for t in shards] if outs is None else list(outs)
    for y, t in zip(outs, shards):
        _check_out(y, t)
    mp = _msrcr_struct(params)
    n = len(shards)
    devs = [t.device.index for t in shards]
    enh = [_tile_enhancer(dv, i if len(set(devs)) == 1 else 0) for i, dv in enumerate(devs)]
    rows = [int(t.shape[0]) for t in shards]
    row0 = [sum(rows[:i]) for i in range(n)]
    for t in shards:  # the library runs on its own streams: order it after the producers of the inputs
        torch.cuda.current_stream(t.device).synchronize()
    VP, I = C.c_void_p * n, C.c_int * n
    _lib.check(_lib.load().llie_msrcr_tiled(VP(*[e._h.value for e in enh]), n, VP(*[t.data_ptr() for t in shards]),
                                            VP(*[y.data_ptr() for y in outs]), I(*row0), I(*rows), sum(rows), w,
                                            C.byref(mp)))
    return outs


def msrcr_tiled_last_ms(shards: Sequence[torch.Tensor]) -> list[float]:
    """Device-side milliseconds each shard's device spent in the last msrcr_tiled() call on these shards."""
    devs = [t.device.index for t in shards]
    ms = []
    for i, dv in enumerate(devs):
        v = C.c_float()
        _lib.check(_lib.load().llie_msrcr_tiled_last_ms(_tile_enhancer(dv, i if len(set(devs)) == 1 else 0)._h, C.byref(v)))
        ms.append(v.value)
    return ms


def _group_enhancers(devices: Sequence[int]) -> list:
    """One Enhancer (ctx) per device slot of a multi-device call; a device listed several times gets one per slot
    ("virtual devices": the multi-device code path on a one-GPU box)."""
    virtual = len(set(devices)) != len(devices)
    return [_tile_enhancer(dv, i if virtual else 0) for i, dv in enumerate(devices)]


def _io_code(dtype) -> int:
    return {torch.float32: 0, torch.uint8: 1, torch.uint16: 2, torch.bfloat16: 3}[dtype]


def enhance_sharded(images, method: str, params: dict | None = None, *, devices: Sequence[int] | None = None, out=None):
    """enhance() over several GPUs of this process (BASELINE.json configs[2] / [3]): independent
    images are split contiguously over `devices`, every device runs the single-GPU sequence on its share, nothing
    is exchanged.  One call into the library (llie_enhance_sharded) drives all devices.

    images : * a NumPy array / CPU tensor (N,H,W,3) or (N,3,H,W): frame i goes to the device that owns it under
               shard_slice(N, d, len(devices)); every device streams its frames host -> device -> host through
               three staging slots on three streams.  Returns one array / tensor of the input's type.
             * a list of CUDA tensors, one per device, already resident on it (zero copy): returns the list of
               results.  `devices` defaults to the tensors' devices.
    A device may be listed several times ("virtual devices").  return_maps is not supported here."""
    params = dict(params or {})
    if method not in _lib.METHOD:
        raise ValueError(f"method must be 'msrcr' or 'zero_dce', got {method!r}")
    lib = _lib.load()
    resident = isinstance(images, (list, tuple))
    if resident:
        ts = list(images)
        if not ts or not all(isinstance(t, torch.Tensor) and t.is_cuda for t in ts):
            raise TypeError("a list of images must hold CUDA tensors, one per device")
        if devices is None:
            devices = [t.device.index for t in ts]
        if len(devices) != len(ts) or any(t.device.index != d for t, d in zip(ts, devices)):
            raise ValueError("tensor i of the list must live on devices[i]")
        first = ts[0]
    else:
        if isinstance(images, torch.Tensor) and images.is_cuda:
            raise ValueError("a single CUDA tensor lives on one device: pass a list with one resident tensor per "
                             "device, or a host array to be split over the devices")
        is_np = isinstance(images, np.ndarray)
        if not is_np and not isinstance(images, torch.Tensor):
            raise TypeError("images must be a numpy.ndarray, a torch.Tensor or a list of CUDA tensors")
        if not devices:
            raise ValueError("devices=[...] is required for host images")
        first = torch.from_numpy(images) if is_np else images
        ts = [first]
    devices = [int(d) for d in devices]
    for t in ts:
        if t.dtype not in _IMAGE_DTYPES or not t.is_contiguous() or t.dim() != 4:
            raise TypeError("images must be contiguous 4-D float32 / uint8 / uint16 (zero_dce: also bfloat16) batches")
        if t.dtype != first.dtype or tuple(t.shape[1:]) != tuple(first.shape[1:]):
            raise ValueError("all shards must share dtype and image shape")
    _, h, w, layout, _ = _resolve_layout(first.shape, params.get("layout"))
    enh = _group_enhancers(devices)
    mp = dp = None
    for e in enh:   # parameter structs are per call; the DCE-Net weights live in every ctx
        mp, dp, want_maps = e._method_structs(method, params)
        if want_maps:
            raise ValueError("enhance_sharded does not return curve maps")
    pstruct = mp if mp is not None else dp
    g = len(devices)
    frame = h * w * 3 * first.element_size()
    if resident:
        outs = [torch.empty_like(t) for t in ts] if out is None else list(out)
        for y, t in zip(outs, ts):
            _check_out(y, t)
            torch.cuda.current_stream(t.device).synchronize()   # the library runs on its own streams
        counts = [int(t.shape[0]) for t in ts]
        ins, outp = [t.data_ptr() for t in ts], [y.data_ptr() for y in outs]
        where = _lib.SHARD_DEVICE
    else:
        y_host = torch.empty_like(first) if out is None else (torch.from_numpy(out) if isinstance(out, np.ndarray) else out)
        _check_out(y_host, first)
        n = int(first.shape[0])
        sl = [shard_slice(n, d, g) for d in range(g)]
        counts = [s_.stop - s_.start for s_ in sl]
        ins = [first.data_ptr() + s_.start * frame for s_ in sl]
        outp = [y_host.data_ptr() + s_.start * frame for s_ in sl]
        where = _lib.SHARD_HOST
    VP, I = C.c_void_p * g, C.c_int * g
    _lib.check(lib.llie_enhance_sharded(VP(*[e._h.value for e in enh]), g, _lib.METHOD[method], _io_code(first.dtype),
                                        VP(*ins), VP(*outp), I(*counts), h, w, layout,
                                        C.cast(C.byref(pstruct), C.c_void_p), where))
    if resident:
        return outs
    return y_host.numpy() if is_np else y_host


def shard_last_ms(devices: Sequence[int]) -> list[float]:
    """Device-side milliseconds of every device's shard in the last enhance_sharded() call on these devices."""
    ms = []
    for e in _group_enhancers([int(d) for d in devices]):
        v = C.c_float()
        _lib.check(_lib.load().llie_shard_last_ms(e._h, C.byref(v)))
        ms.append(v.value)
    return ms


def enhance(images, method: str, params: dict | None = None, *, out=None, device: int | None = None,
            devices: Sequence[int] | None = None):
    """Enhance a low-light image batch on a B200.

    images : numpy.ndarray or torch.Tensor, float32 in [0, 1], uint8 or uint16 (an integer b means b / 255 or
             b / 65535; the result is rint(255 * value) or rint(65535 * value) in the same type, converted inside the
             first / last kernel) -- or, for 'zero_dce' with precision='bf16', torch.bfloat16 tensors in [0, 1]
             (2 bytes per sample on the host link; result rounded to bfloat16) --, shape (H,W,3), (N,H,W,3) or
             (N,3,H,W) (params['layout'] = 'NHWC' | 'NCHW' when the shape is ambiguous).
             CUDA tensors are processed in place on their device and stream (zero copy); NumPy
             arrays and CPU tensors are staged through pinned memory.
    method : 'msrcr'  -- params: sigmas, weights, truncate, alpha, beta, gain, offset, in_scale,
                         in_bias, blur ('pyramid' | 'direct'), finish ('none' | 'minmax' | 'meanstd' | 'percentile': per-image,
                         per-channel stretch of the un-clipped result by min/max, mean +- finish_k std, or the
                         order statistics that clip finish_k percent of the pixels at each end)
             'zero_dce' -- params: weights (79 416-float blob or [(W, b)] * 7), iters, curve_sign,
                         precision ('fp32' = hi/lo-split bf16 on the tensor cores, fp32-accurate |
                         'bf16' | 'fp32_simt' = plain fp32 FMA on the CUDA cores), return_maps
    devices : several GPUs of this process -- the batch is split contiguously over them, no exchange (see
             enhance_sharded(): host batches are streamed through every device, a list of CUDA tensors is
             processed in place, one tensor per device).
    Returns an array/tensor of the input's type, shape and device (and the 24 curve maps when
    return_maps is set).  Raises LlieError on any library failure; there is no CPU fallback.

    Accuracy of the default MSRCR blur ('pyramid'): max-abs <= 1e-4 against the float64 oracle on every named
    config and on full-range content (edges, stripes, noise); where a pure-black pixel (x = 0) has pure-white
    content (x = 1) exactly 4 sigma away the bound is 1.5e-4 (the oracle's hard 4-sigma cut-off cannot be placed
    more finely than the cascade's stride-4 grid; plan_query(sigma) returns the figures).  blur='direct' is the
    exact FIR (3.6e-7) at 1/200 of the speed.
    """
    if devices is not None or isinstance(images, (list, tuple)):
        return enhance_sharded(images, method, params, devices=devices, out=out)
    if isinstance(images, torch.Tensor) and images.is_cuda:
        device = images.device.index
    elif device is None:
        n = C.c_int()
        _lib.check(_lib.load().llie_device_count(C.byref(n)))
        device = torch.cuda.current_device()
    return _enhancer(device).enhance(images, method, params, out=out)


def enhance_stream(frames, method: str, params: dict | None = None, *, depth: int = 3, copy: bool = True,
                   device: int | None = None):
    """Streaming front-end: enhance an iterable of host frames, yielding the results in order.

    frames : iterable of numpy arrays / CPU tensors that all share one shape and dtype -- (H,W,3) frames or small
             (N,H,W,3) / (N,3,H,W) groups, float32 or uint8 -- e.g. a video decoder's output.
    depth  : frames in flight.  Each has its own pinned and device buffers; frame i's upload, frame i-1's kernels
             and frame i-2's download run on three streams, so from depth = 3 the link and the SMs work concurrently.
    copy   : True yields an independent array per frame.  False yields the slot's pinned staging buffer itself (no
             host copy); it is overwritten `depth` frames later, so consume it before advancing that far.
    Frames that are pinned CPU tensors are uploaded straight from the caller's memory (leave them untouched until the
    result has been yielded); anything else is first copied into the slot's pinned stage.
    Yields one result per frame (same type, shape and dtype), delayed by `depth` frames; the tail is flushed when the
    iterable ends.  Same parameters and error behaviour as enhance(); return_maps is not supported.
    """
    if device is None:
        device = torch.cuda.current_device()
    return _enhancer(device).stream(frames, method, params, depth=depth, copy=copy)
