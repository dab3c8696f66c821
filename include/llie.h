/*
 * llie.h -- C ABI of libllie.so: B200-native MSRCR Retinex and Zero-DCE.
 *
 * Drop-in boundary.  The reference (CILAB-IITM/Low_Light_Image_Enhancement) has no
 * code, operator or FFI interface at all (/root/reference/README.md:1-2 is the whole
 * repository), so there is no reference entry point for any function below to
 * "replace".  The boundary is the one BASELINE.json#north_star names -- "Host code is
 * Python calling CUDA through a thin C-ABI (ctypes) layer, exposing one
 * enhance(images, method, params) operator" -- with the concrete contract of
 * SURVEY.md section 8(b).  Each export cites the north_star / SURVEY clause it realises.
 *
 * Conventions: plain C types only; every function returns an int status (0 = OK,
 * negative = LLIE_E_*); nothing throws, aborts or prints; llie_last_error() gives the
 * thread-local text of the last failure.  All image pointers are DEVICE pointers
 * unless the name says "host".  There is NO CPU fallback: a build without CUDA kernels
 * or a box without a B200 fails with LLIE_E_CUDA.
 */
#ifndef LLIE_H_
#define LLIE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LLIE_VERSION 100 /* 0.1.0 */

/* status codes (SURVEY.md 8(b) "errors") */
#define LLIE_OK 0
#define LLIE_E_ARG (-1)
#define LLIE_E_SHAPE (-2)
#define LLIE_E_DTYPE (-3)
#define LLIE_E_CUDA (-4)
#define LLIE_E_NOPEER (-5)
#define LLIE_E_NOMEM (-6)
#define LLIE_E_STATE (-7)

/* layouts of an image batch */
#define LLIE_NHWC 0 /* (N,H,W,3) interleaved RGB */
#define LLIE_NCHW 1 /* (N,3,H,W) planar */

/* element types of Zero-DCE image I/O and arithmetic path */
#define LLIE_F32 0  /* fp32 I/O, fp32-accurate arithmetic on tcgen05: operands split into hi/lo bf16 pairs,
                       a*w = a_hi*w_hi + a_lo*w_hi + a_hi*w_lo, fp32 accumulate (config 1; <= 1e-5 vs fp64) */
#define LLIE_BF16 1 /* fp32 I/O, bf16 operands on tcgen05, fp32 accumulate (config 2) */
#define LLIE_F32_SIMT 2 /* fp32 I/O, plain fp32 FMA on the CUDA cores (GPU-side ground truth, ~6e-8 vs fp64) */

/* MSRCR surround-blur algorithms */
#define LLIE_BLUR_PYRAMID 0 /* decimated Gaussian cascade, O(1) per pixel in sigma; needs sigma >= 12 */
#define LLIE_BLUR_DIRECT 1  /* exact truncated FIR, O(sigma) per pixel; GPU ground truth   */

/* MSRCR global-statistics finishers: applied per image and per channel to the un-clipped
 * value v = (G C R + b) / 255 instead of the plain clip to [0, 1] */
#define LLIE_FINISH_NONE 0    /* out = clip(v, 0, 1)                                                  */
#define LLIE_FINISH_MINMAX 1  /* out = (v - min) / (max - min)                                        */
#define LLIE_FINISH_MEANSTD 2 /* out = clip((v - (mean - k std)) / (2 k std), 0, 1), population std */
#define LLIE_FINISH_PERCENTILE 3 /* out = clip((v - lo) / (hi - lo), 0, 1); lo / hi = the r-th smallest / largest
                                    value, r = floor(N k / 100) ("simplest colour balance", k percent per tail) */

#define LLIE_MAX_SCALES 8

/* north_star (a): "multi-scale Gaussian surround at sigma=15/80/250, log-ratio, colour
 * restoration and gain/offset"; defaults pinned by SURVEY.md 9.1. */
typedef struct llie_msrcr_params {
  uint32_t struct_size; /* = sizeof(llie_msrcr_params) */
  int32_t n_scales;
  float sigma[LLIE_MAX_SCALES];
  float weight[LLIE_MAX_SCALES];
  float truncate; /* kernel radius = ceil(truncate * sigma) */
  float alpha, beta, gain, offset;
  float in_scale, in_bias; /* I = in_scale * x + in_bias */
  int32_t blur;            /* LLIE_BLUR_* */
  int32_t finish;          /* LLIE_FINISH_* (fp32 images, LLIE_BLUR_PYRAMID) */
  float finish_k;          /* k of LLIE_FINISH_MEANSTD (default 2) / percent per tail of LLIE_FINISH_PERCENTILE */
} llie_msrcr_params;

/* north_star (b): "8 iterations of LE(x)=x+alpha*x*(1-x)"; SURVEY.md 9.2. */
typedef struct llie_dce_params {
  uint32_t struct_size; /* = sizeof(llie_dce_params) */
  int32_t iters;        /* curve iterations; the net always emits 3*8 maps, iters <= 8 */
  int32_t curve_sign;   /* +1: x + a x (1-x) (north_star); -1: upstream convention */
  int32_t precision;    /* LLIE_F32, LLIE_BF16 or LLIE_F32_SIMT */
} llie_dce_params;

typedef struct llie_ctx llie_ctx;

/* library identity / diagnostics */
int llie_version(void);
const char* llie_last_error(void);
/* fills *count with the number of CUDA devices; LLIE_E_CUDA if none (no CPU fallback) */
int llie_device_count(int* count);

/* Fill a params struct with the SURVEY.md section 9 defaults.  The structs are versioned by their leading
 * struct_size: the CALLER sets p->struct_size = sizeof(its struct) BEFORE the call and the library writes exactly
 * that many bytes (LLIE_E_ARG for 0 or any size it does not know).  llie_msrcr_params is accepted in its first
 * layout (104 bytes, through `blur`) and the current one (112 bytes); a shorter struct means finish = NONE. */
int llie_msrcr_default_params(llie_msrcr_params* p);
int llie_dce_default_params(llie_dce_params* p);

/* A context binds one device (SURVEY.md 8(b) "ownership": streams, workspaces and
 * packed weights belong to the ctx; no allocation on the hot path after warm-up).  A ctx is not thread-safe: one
 * ctx per host thread; different ctxs share no state and may be driven from different threads concurrently. */
int llie_create(llie_ctx** out, int device);
int llie_destroy(llie_ctx* ctx);
/* SURVEY.md 8(b) `llie_create(out, device_ids, n_dev)`: one ctx per listed device, all or nothing (out_ctxs has
 * n_dev slots).  The same device may be listed more than once ("virtual devices" for one-GPU tests). */
int llie_create_group(llie_ctx** out_ctxs, const int* device_ids, int n_dev);
int llie_destroy_group(llie_ctx* const* ctxs, int n_dev);
int llie_sync(llie_ctx* ctx);
/* bytes of device workspace currently held by the ctx */
int llie_workspace_bytes(llie_ctx* ctx, size_t* bytes);
/* kernels launched by this ctx since creation (bench.py's gpu_launches) */
int llie_launch_count(llie_ctx* ctx, uint64_t* count);

/* Per-kernel device timing for bench.py's roofline figure (SURVEY.md 8(d) "evidence"): between
 * begin and end every kernel this ctx launches is bracketed by CUDA events on its launching
 * stream.  end synchronises and writes "kernel_name:launches:total_ms;" records into buf. */
int llie_profile_begin(llie_ctx* ctx);
int llie_profile_end(llie_ctx* ctx, char* buf, size_t buflen);

/* north_star kernel (1): MSRCR on a device batch. in/out: n*h*w*3 fp32 in `layout`,
 * values in [0,1].  stream: cudaStream_t as void*; NULL = the ctx's own stream and a
 * synchronous return.  out may not alias in. */
int llie_msrcr_f32(llie_ctx* ctx, const float* in, float* out, int n, int h, int w, int layout,
                   const llie_msrcr_params* params, void* stream);

/* MSRCR of ONE h x w x 3 fp32 image without a ctx and without any allocation -- NO COUNTERPART IN THE REFERENCE (nor
 * in SURVEY.md 8(b)); it exists for callers that cannot let this library call cudaMalloc: a plug-in inside an
 * application that has already claimed the whole GPU for its own pool (the library's cudaMalloc would fail there),
 * a frame loop that cannot afford the device-wide synchronisation a cudaMalloc / cudaFree implies, or a runtime whose
 * policy forbids device allocation after start-up.  in / out are device pointers on the CURRENT device (out may not
 * alias in); everything in between lives in one fixed array of LLIE_STATIC_SCRATCH_BYTES that is part of the library's
 * device image, so every GPU that uses this function gives up that much memory until the process exits.
 *   - Arithmetic: the exact truncated FIR (what llie_msrcr_f32 does for blur = LLIE_BLUR_DIRECT; params->blur is
 *     ignored) and the plain clip; params->finish must be LLIE_FINISH_NONE.
 *   - Limits: (1 + n_scales) * h * w * 12 bytes + 4 * LLIE_STATIC_MAX_TAPS must fit the array (three scales: 697 685
 *     pixels, e.g. 960 x 540), and the filters together may have at most LLIE_STATIC_MAX_TAPS taps (each scale uses
 *     2 * ceil(truncate * sigma) + 1, rounded up to a multiple of 4).  Anything larger: LLIE_E_NOMEM, nothing is run.
 *   - The first call on a device looks up the array's address there (and makes the driver load this part of the
 *     library onto that GPU); make that call during start-up, before any stream capture is begun.
 *   - All calls on one GPU share the array from its first byte.  Calls that are ordered on the device (same stream,
 *     or events between streams) are fine; calls that can overlap on the GPU corrupt each other silently.  The array
 *     keeps the last call's intermediates; the function needs none of it cleared.
 *   - stream: cudaStream_t as void*, work is enqueued and the call returns; NULL = the default stream and a
 *     synchronous return.  No cudaMalloc, cudaFree or other allocating call is made, on any call. */
#define LLIE_STATIC_SCRATCH_BYTES (32u << 20) /* 32 MiB per GPU that loads it */
#define LLIE_STATIC_MAX_TAPS 16384
int llie_msrcr_direct_static_f32(const float* in, float* out, int h, int w, int layout,
                                 const llie_msrcr_params* params, void* stream);

/* SURVEY.md 8(f) rank 1 -- 8-bit images: in/out are n*h*w*3 uint8 device arrays in `layout`; a byte b means the
 * value b/255 and the result is stored as rint(255 * value).  The conversions happen inside the first and last
 * kernels of the path (no extra pass over the image), so the image costs 1 byte per sample on HBM and on PCIe.
 * LLIE_BLUR_PYRAMID only. */
int llie_msrcr_u8(llie_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w, int layout,
                  const llie_msrcr_params* params, void* stream);

/* 16-bit images (the other half of 8(f) rank 1): as llie_msrcr_u8 with a sample b meaning b/65535 and the result
 * stored as rint(65535 * value); 2 bytes per sample on HBM and on PCIe. */
int llie_msrcr_u16(llie_ctx* ctx, const uint16_t* in, uint16_t* out, int n, int h, int w, int layout,
                   const llie_msrcr_params* params, void* stream);

/* Host-side planner of the pyramid blur, exposed so that CPU tests can check it without a
 * GPU: for `sigma`/`truncate` returns the cascade level's stride (4 or 16), the radius of the
 * low-resolution filter and the square root of its second moment (its taps are free-form,
 * fitted by weighted least squares), and the L1 distance between the cascade's effective kernel
 * and the oracle's truncated Gaussian.  Needs no device. */
int llie_msrcr_plan_query(float sigma, float truncate, int* stride, double* sigma_low, int* radius_low,
                          double* l1_error);

/* Accuracy figures of that plan, all for the default map I = 255 x + 1 (a black pixel is I = 1, a white one 256):
 *   l1_error    sum |k_cascade - k_oracle|, phase average: bounds the relative error of a surround on fine texture;
 *   edge_error  largest relative error |dS| / S of the surround of a BLACK pixel beside a WHITE half-plane, over
 *               all distances from the edge (the oracle's hard 4-sigma cut-off cannot be placed more finely than
 *               the stride-4 grid);
 *   line_error  the same for a 1-pixel white line on black.
 * The MSRCR output moves by (gain C / 255) weight_k * (that figure) -- 0.135 x with the defaults (C = 171.6 for a
 * grey-black pixel) -- so the 1e-4 parity budget of the fp32 path holds on such content only while the figure is
 * below 7.4e-4; see DESIGN.md 2.1 for the measured bounds per content class.  LLIE_BLUR_DIRECT has no such term. */
int llie_msrcr_plan_errors(float sigma, float truncate, double* l1_error, double* edge_error, double* line_error);

/* The low-resolution taps of that plan (2 radius_low + 1 floats, centred).  *count receives the number of taps;
 * taps may be NULL with capacity 0 to query the count only. */
int llie_msrcr_plan_taps(float sigma, float truncate, float* taps, int capacity, int* count);

/* north_star kernel (2): upload DCE-Net weights.  host_packed: the 79 416-float blob
 * W1 b1 ... W7 b7 (OIHW) of SURVEY.md 9.2, in HOST memory. */
int llie_dce_set_weights(llie_ctx* ctx, const float* host_packed, size_t count);

/* north_star kernels (2)+(3): Zero-DCE forward.  in/out: n*h*w*3 fp32 device arrays in
 * `layout`; maps_or_null: optional n*h*w*24 fp32 output of the tanh curve maps (same
 * layout family: NHWC -> (N,H,W,24), NCHW -> (N,24,H,W)). */
int llie_dce_forward(llie_ctx* ctx, const float* in, float* out, float* maps_or_null, int n, int h, int w,
                     int layout, const llie_dce_params* params, void* stream);

/* 8-bit variant of llie_dce_forward (same conventions as llie_msrcr_u8; LLIE_BF16 precision only; the curve
 * maps, if requested, stay fp32). */
int llie_dce_forward_u8(llie_ctx* ctx, const uint8_t* in, uint8_t* out, float* maps_or_null, int n, int h, int w,
                        int layout, const llie_dce_params* params, void* stream);

/* 16-bit variant (same conventions as llie_msrcr_u16; LLIE_BF16 precision only). */
int llie_dce_forward_u16(llie_ctx* ctx, const uint16_t* in, uint16_t* out, float* maps_or_null, int n, int h, int w,
                         int layout, const llie_dce_params* params, void* stream);

/* bfloat16 images ("bf16 torch tensors allowed for zero_dce bf16"): in / out are n*h*w*3 bfloat16
 * values in [0, 1] (the bit layout of __nv_bfloat16 / torch.bfloat16), read by the first kernel and rounded to nearest
 * even by the last one, so the image costs 2 bytes per sample on HBM and on PCIe.  LLIE_BF16 precision only; the curve
 * maps, if requested, stay fp32.  Internally nothing changes: the input is rounded to bf16 for the tensor cores anyway,
 * so the only extra error against llie_dce_forward is the rounding of the stored result (<= 2^-9 relative). */
int llie_dce_forward_bf16io(llie_ctx* ctx, const void* in, void* out, float* maps_or_null, int n, int h, int w,
                            int layout, const llie_dce_params* params, void* stream);

/* Test hook: after a LLIE_BF16 or LLIE_F32 forward, copies inter-layer activation `which` (1..6 = x1..x6 of
 * SURVEY.md 9.2) of frame `frame` of the last processed chunk to HOST memory as fp32 [32][h][w]. */
int llie_debug_dce_activation(llie_ctx* ctx, int which, int frame, float* host_out);

/* Test hook for the LLIE_BF16 path: bit 0 = also write the activations that normally never leave the
 * chip (x4, x6) so that llie_debug_dce_activation can read them; bit 1 = run one launch per layer
 * instead of the fused conv1-3 / conv4-5 / conv6-7 chains (same kernel, for A/B measurements). */
int llie_debug_dce_mode(llie_ctx* ctx, int mode);

/* north_star kernel (3) alone: curves from given maps (N,H,W,24)/(N,24,H,W) fp32. */
int llie_dce_curves(llie_ctx* ctx, const float* in, const float* maps, float* out, int n, int h, int w,
                    int layout, const llie_dce_params* params, void* stream);

/* Batch sharder (north_star: "shards batches of independent images or frames across the
 * 8xB200 box with no collectives").  ONE call processes n_dev independent sub-batches: ctxs[d] enhances the
 * n_per_dev[d] images at in_per_dev[d] into out_per_dev[d] (h x w x 3 each, `layout`, samples of io_type) with the
 * unchanged single-device launch sequence on its own stream.  Nothing is exchanged between devices; every device
 * is synchronised once, at the end (the call is synchronous).  params: const llie_msrcr_params* or
 * const llie_dce_params* according to method (Zero-DCE: llie_dce_set_weights on every ctx first; no curve maps).
 * where = LLIE_SHARD_DEVICE: the pointers are device memory on ctxs[d]'s device (zero copy).
 * where = LLIE_SHARD_HOST:   the pointers are HOST memory; each device streams its frames through three staging
 *                            slots on three streams (upload / kernels / download overlap).  Pinned memory is
 *                            copied in place, pageable memory through pinned stages owned by the ctx.
 * The ctxs must be distinct and live on n_dev distinct devices, or all on one device ("virtual devices"). */
#define LLIE_METHOD_MSRCR 0
#define LLIE_METHOD_ZERO_DCE 1
#define LLIE_IO_F32 0
#define LLIE_IO_U8 1
#define LLIE_IO_U16 2
#define LLIE_IO_BF16 3 /* Zero-DCE (LLIE_BF16 precision) only */
#define LLIE_SHARD_DEVICE 0
#define LLIE_SHARD_HOST 1
int llie_enhance_sharded(llie_ctx* const* ctxs, int n_dev, int method, int io_type, const void* const* in_per_dev,
                         void* const* out_per_dev, const int* n_per_dev, int h, int w, int layout, const void* params,
                         int where);
/* Device-side duration (CUDA events on the ctx's stream) of this ctx's shard in the last llie_enhance_sharded call. */
int llie_shard_last_ms(llie_ctx* ctx, float* ms);

/* Row-tiled MSRCR of ONE image across several devices of this process (config 5):
 * device d holds rows [row0[d], row0[d]+rows[d]) of the h_total x w image in
 * in_shard[d]/out_shard[d] (NHWC fp32).  Low-resolution halo rows are exchanged with
 * cudaMemcpyPeerAsync over NVLink (no NCCL).  ctxs[d] must live on distinct devices with
 * peer access, or all on one device ("virtual devices" for single-GPU tests). */
int llie_msrcr_tiled(llie_ctx* const* ctxs, int n_dev, const float* const* in_shard, float* const* out_shard,
                     const int* row0, const int* rows, int h_total, int w, const llie_msrcr_params* params);

/* Device-side duration (CUDA events on the shard's stream, first copy to last kernel) of this ctx's
 * shard in the last successful llie_msrcr_tiled call.  All shards start after a common host barrier,
 * so the max over the ctxs is the multi-GPU time of the call (SURVEY.md 8(d) "timing method"). */
int llie_msrcr_tiled_last_ms(llie_ctx* ctx, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* LLIE_H_ */
