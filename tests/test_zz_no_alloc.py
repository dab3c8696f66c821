"""llie_msrcr_direct_static_f32 -- the one entry point of libllie that never allocates.

It is meant for hosts where this library must not (or cannot) call cudaMalloc: a plug-in inside an application that
has already taken the whole GPU for its own pool, a frame loop that cannot take the device-wide stall of an
allocation, a runtime that forbids allocating after start-up.  Its intermediates live in a fixed 32 MiB array inside
the library's device image, shared by all calls on a GPU and never cleared.

What is checked here, through the C ABI (ctypes on the in-tree libllie.so):
  * two calls in a row (different images, shapes and layouts, the second on a side stream, the output pre-filled
    with NaN) each reproduce the existing entry point llie_msrcr_f32(blur = DIRECT) -- so leftovers of the first
    call in the shared array do not leak into the second -- and both meet the fp32 parity budget against the oracle;
  * free device memory does not move across further calls (nothing is allocated or freed);
  * problems beyond the fixed array are refused with LLIE_E_NOMEM and a message, before any CUDA work (needs no GPU).

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure in this out-of-scope extra
must not hide the parity tests.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest
import torch

import oracle
import low_light_image_enhancement_b200 as L
from low_light_image_enhancement_b200 import _lib, api

TOL_F32 = 1e-4      # the fp32 parity budget of tests/test_gpu_parity.py (north_star: max-abs on [0, 1] pixels)
TOL_SAME = 2e-6     # new vs existing entry point: same kernels, taps rounded from double on device vs on host


def _static_call(lib, x: torch.Tensor, out: torch.Tensor, h, w, layout, params, stream_ptr=None):
    return lib.llie_msrcr_direct_static_f32(x.data_ptr(), out.data_ptr(), h, w, layout, C.byref(params), stream_ptr)


@pytest.mark.gpu
def test_static_scratch_matches_existing_entry_point_twice(lib_built):
    lib = lib_built
    p = api._msrcr_struct(dict(blur="direct"))
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)

    # call 1: NHWC, default (NULL) stream = synchronous return; fills most of the array's first megabytes
    a = oracle.synth_lowlight(300, 420, seed=11)
    xa = torch.from_numpy(a).to(dev)
    ya = torch.full_like(xa, float("nan"))
    _lib.check(_static_call(lib, xa, ya, 300, 420, _lib.NHWC, p))
    ref_a = L.enhance(xa, "msrcr", dict(blur="direct"))
    err_a = float((ya - ref_a).abs().max())
    assert torch.isfinite(ya).all() and err_a <= TOL_SAME, err_a
    assert np.abs(ya.cpu().numpy() - oracle.msrcr(a)).max() <= TOL_F32

    # call 2 straight after: smaller image, planar layout, other scales, on a side stream ordered after call 1
    b = oracle.synth_lowlight(97, 131, seed=12)
    q = dict(sigmas=(12.0, 33.0), weights=(0.4, 0.6), gain=0.8, offset=120.0, blur="direct")
    pq = api._msrcr_struct(q)
    xb = torch.from_numpy(np.ascontiguousarray(b.transpose(2, 0, 1))).to(dev)
    yb = torch.full_like(xb, float("nan"))
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    _lib.check(_static_call(lib, xb, yb, 97, 131, _lib.NCHW, pq, C.c_void_p(side.cuda_stream)))
    side.synchronize()
    ref_b = L.enhance(xb[None], "msrcr", dict(q, layout="NCHW"))[0]
    err_b = float((yb - ref_b).abs().max())
    assert torch.isfinite(yb).all() and err_b <= TOL_SAME, err_b
    ob = oracle.msrcr(b, {k: v for k, v in q.items() if k != "blur"})
    assert np.abs(yb.cpu().numpy().transpose(1, 2, 0) - ob).max() <= TOL_F32

    # the first image again: identical bits to call 1 although the array now holds call 2's leftovers
    ya2 = torch.full_like(xa, float("nan"))
    _lib.check(_static_call(lib, xa, ya2, 300, 420, _lib.NHWC, p))
    assert torch.equal(ya2, ya)

    # nothing is allocated or freed by further calls: the driver's free-memory figure stays put
    torch.cuda.synchronize(dev)
    free0, _ = torch.cuda.mem_get_info(dev)
    for _ in range(3):
        _lib.check(_static_call(lib, xa, ya2, 300, 420, _lib.NHWC, p))
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0, (free0, free1)


def test_static_scratch_refuses_what_does_not_fit(lib_built):
    """Limits are checked on the host before anything touches CUDA, so this runs without a GPU; the pointers are
    never dereferenced."""
    lib = lib_built
    p = api._msrcr_struct(dict(blur="direct"))
    fake_in, fake_out = C.c_void_p(0x1000), C.c_void_p(0x2000)
    call = lib.llie_msrcr_direct_static_f32
    # 1080p with three scales needs 4 planes of 24.9 MB: far beyond 32 MiB
    assert call(fake_in, fake_out, 1080, 1920, _lib.NHWC, C.byref(p), None) == _lib.E_NOMEM
    msg = lib.llie_last_error().decode()
    assert "static scratch" in msg and str(_lib.STATIC_SCRATCH_BYTES) in msg
    # the largest 3-scale image that fits is 697 685 pixels: one more row of a 835-wide image tips it over
    fits = (_lib.STATIC_SCRATCH_BYTES // 4 - _lib.STATIC_MAX_TAPS) // (3 * 4)
    assert fits == 697685
    assert call(fake_in, fake_out, fits // 835 + 1, 835, _lib.NHWC, C.byref(p), None) == _lib.E_NOMEM
    # filters longer than the tap area
    big = api._msrcr_struct(dict(sigmas=(2500.0,), weights=(1.0,), blur="direct"))
    assert call(fake_in, fake_out, 16, 16, _lib.NHWC, C.byref(big), None) == _lib.E_NOMEM
    assert "taps" in lib.llie_last_error().decode()
    # finishers, aliasing and bad shapes are argument errors
    fin = api._msrcr_struct(dict(finish="minmax"))
    assert call(fake_in, fake_out, 16, 16, _lib.NHWC, C.byref(fin), None) == _lib.E_ARG
    assert call(fake_in, fake_in, 16, 16, _lib.NHWC, C.byref(p), None) == _lib.E_ARG
    assert call(fake_in, fake_out, 0, 16, _lib.NHWC, C.byref(p), None) == _lib.E_SHAPE
