// llie_msrcr_direct_static_f32: exact-FIR MSRCR of ONE image that allocates nothing.
//
// Who needs this.  Every other entry point of libllie takes a ctx, and a ctx grows device workspaces with cudaMalloc
// the first time it sees a shape.  That is unusable for
//   * a plug-in living inside a host application that reserved the whole GPU for its own memory pool at start-up
//     (a cudaMalloc from the plug-in then simply fails with out-of-memory, however small the request);
//   * a frame loop with a latency budget: cudaMalloc / cudaFree synchronise the whole device, so one call from this
//     library would stall every other stream of the process for the duration;
//   * a caller whose policy (a certified or sandboxed runtime) forbids device allocation after initialisation.
// Such callers hand in their own input and output buffers and this file supplies the intermediate storage itself:
// one fixed array in the library's device image, LLIE_STATIC_SCRATCH_BYTES large, which exists on every GPU that
// loads this module from the moment it is loaded until the process exits.  The price is the fixed maximum problem
// size (checked, LLIE_E_NOMEM beyond it) and the loss of re-entrancy described below.
//
// The arithmetic is that of msrcr_direct.cu (LLIE_BLUR_DIRECT): the same fir_x / fir_y kernels from fir.cuh and the
// same per-pixel formula, so its results agree with llie_msrcr_f32(blur = DIRECT) to rounding of the filter taps
// (here they are produced by a kernel, there on the host; both in double, rounded once to fp32).
//
// Sharing.  There is ONE array per GPU and every call on that GPU uses it from offset 0, whatever the stream or the
// calling thread.  Two calls whose kernels can overlap on the device (different streams without an ordering between
// them) overwrite each other's intermediates and produce garbage without any error.  Calls issued to one stream, or
// otherwise ordered, are safe.  The array keeps whatever the previous call left in it; nothing in the algorithm
// relies on a zeroed start (taps, row-filtered plane and surrounds are each fully written before they are read, and
// the tail of every tap block is cleared by the tap kernel), so no clearing pass is needed.
//
// No counterpart in the reference (which has no code at all: /root/reference/README.md:1-2).
#include "fir.cuh"

#include <cmath>
#include <mutex>

namespace llie {
namespace {

// The scratch itself.  In HOST code this name is only a handle for cudaGetSymbolAddress -- it is not a device
// address and must never be passed to a kernel; static_scratch_for_current_device() below resolves it.
static __device__ __align__(256) unsigned char g_static_scratch[LLIE_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
float* g_scratch_addr[kMaxDevices] = {};  // per device: the array's device address, resolved once

// First call on a device: resolves the array's address there (this also makes the driver load this module's device
// image, which is when the array comes into existence on that GPU).  Later calls: a table look-up.
int static_scratch_for_current_device(float** out) {
  int dev = -1;
  LLIE_CUDA_OK(cudaGetDevice(&dev));
  LLIE_REQUIRE(dev >= 0 && dev < kMaxDevices, LLIE_E_ARG, "device ordinal %d is beyond the %d this entry point tracks", dev,
               kMaxDevices);
  std::lock_guard<std::mutex> lk(g_setup_mu);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    LLIE_CUDA_OK(cudaGetSymbolAddress(&p, g_static_scratch));
    LLIE_REQUIRE(p != nullptr, LLIE_E_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
    g_scratch_addr[dev] = static_cast<float*>(p);
  }
  *out = g_scratch_addr[dev];
  return LLIE_OK;
}

struct TapPlan {
  int n_scales;
  float sigma[LLIE_MAX_SCALES];
  int radius[LLIE_MAX_SCALES];
  int offset[LLIE_MAX_SCALES];  // first tap of scale s, in floats from the start of the tap area
  int padded[LLIE_MAX_SCALES];  // floats reserved for scale s (2 r + 1 rounded up to a multiple of 4)
};

// Block s writes the normalised truncated Gaussian of scale s: k[t] = exp(-t^2 / (2 sigma^2)) / sum, |t| <= r,
// evaluated and normalised in double and rounded once to fp32 (as msrcr_direct_run does on the host), then clears
// the padding behind it.
__global__ void __launch_bounds__(256) static_taps_kernel(float* __restrict__ taps, TapPlan tp) {
  __shared__ double part[256];
  const int s = blockIdx.x, r = tp.radius[s], len = 2 * r + 1;
  const double sg = (double)tp.sigma[s], den = 2.0 * sg * sg;
  double acc = 0.0;
  for (int i = threadIdx.x; i < len; i += blockDim.x) {
    const double t = (double)(i - r);
    acc += exp(-t * t / den);
  }
  part[threadIdx.x] = acc;
  __syncthreads();
  for (int half = 128; half > 0; half >>= 1) {
    if ((int)threadIdx.x < half) part[threadIdx.x] += part[threadIdx.x + half];
    __syncthreads();
  }
  const double sum = part[0];
  float* dst = taps + tp.offset[s];
  for (int i = threadIdx.x; i < tp.padded[s]; i += blockDim.x) {
    const double t = (double)(i - r);
    dst[i] = i < len ? (float)(exp(-t * t / den) / sum) : 0.f;
  }
}

struct StaticEpi {
  int n_scales;
  float weight[LLIE_MAX_SCALES];
  float alpha, beta, gain, offset, in_scale, in_bias;
};

// The per-pixel tail of MSRCR with the formula and operation order of msrcr_direct.cu's epilogue:
// R_c = sum_s w_s (ln I_c - ln S_{s,c}); C_c = beta (ln(alpha I_c) - ln sum_c I_c); out = clip((G C R + b) / 255).
// The surrounds of scale s start at surround + s * scale_stride and have the image's own layout.
__global__ void static_msrcr_epilogue_kernel(const float* __restrict__ x, const float* __restrict__ surround,
                                             long long scale_stride, float* __restrict__ out, long long n_px, ImgView v,
                                             int W, StaticEpi p) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n_px) return;
  const long long yy = i / W, xx = i - yy * W;
  const long long base = yy * v.sy + xx * v.sx;
  float I[3], lI[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    I[c] = fmaf(p.in_scale, x[base + c * v.sc], p.in_bias);
    lI[c] = logf(I[c]);
  }
  const float lsum = logf(I[0] + I[1] + I[2]);
  const float lalpha = logf(p.alpha);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float R = 0.f;
    for (int s = 0; s < p.n_scales; ++s)
      R = fmaf(p.weight[s], lI[c] - logf(surround[s * scale_stride + base + c * v.sc]), R);
    const float Cc = p.beta * (lalpha + lI[c] - lsum);
    const float o = fmaf(p.gain * Cc, R, p.offset) * (1.0f / 255.0f);
    out[base + c * v.sc] = fminf(fmaxf(o, 0.f), 1.f);
  }
}

#define LLIE_STATIC_LAUNCH_OK(what)                                                          \
  do {                                                                                       \
    cudaError_t _e = cudaGetLastError();                                                     \
    LLIE_REQUIRE(_e == cudaSuccess, LLIE_E_CUDA, "launch of %s failed: %s", what, cudaGetErrorString(_e)); \
  } while (0)

}  // namespace
}  // namespace llie

using namespace llie;

extern "C" int llie_msrcr_direct_static_f32(const float* in, float* out, int h, int w, int layout,
                                            const llie_msrcr_params* params, void* stream) {
  LLIE_REQUIRE(in && out, LLIE_E_ARG, "null image pointer");
  LLIE_REQUIRE(in != out, LLIE_E_ARG, "in-place operation is not supported (out aliases in)");
  LLIE_REQUIRE(h > 0 && w > 0, LLIE_E_SHAPE, "bad shape h=%d w=%d", h, w);
  LLIE_REQUIRE(layout == LLIE_NHWC || layout == LLIE_NCHW, LLIE_E_ARG, "unknown layout %d", layout);
  llie_msrcr_params p;
  LLIE_TRY(msrcr_params_normalise(params, &p));
  LLIE_REQUIRE(p.n_scales >= 1 && p.n_scales <= LLIE_MAX_SCALES, LLIE_E_ARG, "n_scales %d out of range", p.n_scales);
  for (int s = 0; s < p.n_scales; ++s)
    LLIE_REQUIRE(p.sigma[s] > 0.f && p.sigma[s] <= 4096.f, LLIE_E_ARG, "sigma[%d]=%g out of range", s, p.sigma[s]);
  LLIE_REQUIRE(p.truncate > 0.f && p.truncate <= 8.f, LLIE_E_ARG, "truncate %g out of range", p.truncate);
  LLIE_REQUIRE(p.in_bias >= 1e-30f && p.in_bias <= 1e30f && p.in_scale >= 0.f && p.in_scale <= 1e30f, LLIE_E_ARG,
               "in_bias must lie in [1e-30, 1e30] and in_scale in [0, 1e30] (got %g, %g)", p.in_bias, p.in_scale);
  LLIE_REQUIRE(p.finish == LLIE_FINISH_NONE, LLIE_E_ARG,
               "the static-scratch entry point has no global-statistics finishers (finish must be LLIE_FINISH_NONE)");

  // ---- does the problem fit the fixed array?  [ taps | row-filtered plane | surround of scale 0 | 1 | ... ] ----
  TapPlan tp;
  tp.n_scales = p.n_scales;
  int n_taps = 0, rmax = 0;
  for (int s = 0; s < p.n_scales; ++s) {
    const int r = (int)std::ceil((double)p.truncate * (double)p.sigma[s]);
    tp.sigma[s] = p.sigma[s];
    tp.radius[s] = r;
    tp.offset[s] = n_taps;
    tp.padded[s] = (2 * r + 1 + 3) & ~3;
    n_taps += tp.padded[s];
    rmax = r > rmax ? r : rmax;
  }
  LLIE_REQUIRE(n_taps <= LLIE_STATIC_MAX_TAPS, LLIE_E_NOMEM,
               "static scratch: the %d scales need %d filter taps, the fixed tap area holds %d (use smaller sigma / truncate, "
               "or llie_msrcr_f32 with a ctx)", p.n_scales, n_taps, LLIE_STATIC_MAX_TAPS);
  const size_t img_elems = (size_t)h * w * 3;
  const size_t need = ((size_t)LLIE_STATIC_MAX_TAPS + (size_t)(1 + p.n_scales) * img_elems) * sizeof(float);
  LLIE_REQUIRE(need <= (size_t)LLIE_STATIC_SCRATCH_BYTES, LLIE_E_NOMEM,
               "static scratch: %d x %d with %d scales needs %zu bytes of intermediates, the fixed array holds %zu "
               "(at most %zu pixels at this number of scales; larger images need llie_msrcr_f32 with a ctx)",
               h, w, p.n_scales, need, (size_t)LLIE_STATIC_SCRATCH_BYTES,
               ((size_t)LLIE_STATIC_SCRATCH_BYTES / sizeof(float) - LLIE_STATIC_MAX_TAPS) / (3 * (size_t)(1 + p.n_scales)));
  const int C = (layout == LLIE_NHWC) ? 3 : 1;
  const size_t smem_max = (size_t)(128 + 2 * rmax + 3) * C * sizeof(float);
  LLIE_REQUIRE(smem_max <= 200 * 1024, LLIE_E_ARG, "direct blur: radius %d too large for shared-memory staging", rmax);

  float* scratch = nullptr;
  LLIE_TRY(static_scratch_for_current_device(&scratch));
  if (smem_max > 48 * 1024)
    LLIE_CUDA_OK(cudaFuncSetAttribute(fir_x_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max));

  cudaStream_t st = (cudaStream_t)stream;
  float* d_taps = scratch;
  float* d_tmp = scratch + LLIE_STATIC_MAX_TAPS;
  float* d_sur = d_tmp + img_elems;

  static_taps_kernel<<<p.n_scales, 256, 0, st>>>(d_taps, tp);
  LLIE_STATIC_LAUNCH_OK("static_taps_kernel");

  const long long rows = (long long)h * (3 / C);  // NHWC: h rows of 3 w floats; NCHW: 3 h rows of w floats
  const int planes = 3 / C, RW = w * C;
  for (int s = 0; s < p.n_scales; ++s) {
    const float* k = d_taps + tp.offset[s];
    const int r = tp.radius[s];
    const dim3 gx((unsigned)rows, (w + 127) / 128);
    const size_t sm = (size_t)(128 + 2 * r + 3) * C * sizeof(float);
    fir_x_kernel<true><<<gx, 128, sm, st>>>(in, d_tmp, k, r, w, C, p.in_scale, p.in_bias, (int)rows, 0, 0);
    LLIE_STATIC_LAUNCH_OK("fir_x_kernel");
    const dim3 gy((RW + 127) / 128, (h + 3) / 4, planes);
    fir_y_kernel<<<gy, 128, 0, st>>>(d_tmp, d_sur + (size_t)s * img_elems, k, r, h, RW, 0, h, 0, h, 0, h);
    LLIE_STATIC_LAUNCH_OK("fir_y_kernel");
  }
  StaticEpi ep;
  ep.n_scales = p.n_scales;
  for (int s = 0; s < p.n_scales; ++s) ep.weight[s] = p.weight[s];
  ep.alpha = p.alpha; ep.beta = p.beta; ep.gain = p.gain; ep.offset = p.offset;
  ep.in_scale = p.in_scale; ep.in_bias = p.in_bias;
  const long long npx = (long long)h * w;
  static_msrcr_epilogue_kernel<<<(unsigned)((npx + 255) / 256), 256, 0, st>>>(in, d_sur, (long long)img_elems, out, npx,
                                                                           make_view(layout, 3, h, w), w, ep);
  LLIE_STATIC_LAUNCH_OK("static_msrcr_epilogue_kernel");
  if (!stream) LLIE_CUDA_OK(cudaStreamSynchronize(st));
  return LLIE_OK;
}
