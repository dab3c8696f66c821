// C-ABI of libllie.so (include/llie.h): contexts, argument validation and dispatch.
// int status codes, thread-local error text, no exceptions across the boundary,
// no CPU fallback -- every entry point that computes needs a CUDA device.
#include "llie_common.cuh"

#include <cstring>
#include <new>

namespace llie {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

const char* last_error_text() { return g_err; }

// RAII device switch so that a ctx can be driven from a thread whose current device differs.
struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

static int check_batch(const void* in, const void* out, int n, int h, int w, int layout) {
  LLIE_REQUIRE(in && out, LLIE_E_ARG, "null image pointer");
  LLIE_REQUIRE(in != out, LLIE_E_ARG, "in-place operation is not supported (out aliases in)");
  LLIE_REQUIRE(n >= 0 && h > 0 && w > 0, LLIE_E_SHAPE, "bad shape n=%d h=%d w=%d", n, h, w);
  LLIE_REQUIRE((long long)h * w <= (1ll << 31), LLIE_E_SHAPE, "image too large: %d x %d", h, w);
  LLIE_REQUIRE(layout == LLIE_NHWC || layout == LLIE_NCHW, LLIE_E_ARG, "unknown layout %d", layout);
  return LLIE_OK;
}

}  // namespace llie

using namespace llie;

int llie_ctx::workspace(const char* name, size_t bytes, void** out) {
  auto it = bufs.find(name);
  if (it != bufs.end() && it->second.second >= bytes) {
    *out = it->second.first;
    return LLIE_OK;
  }
  if (it != bufs.end()) {
    // the old buffer may still be in use by queued work
    LLIE_CUDA_OK(cudaDeviceSynchronize());
    cudaFree(it->second.first);
    bufs.erase(it);
  }
  void* p = nullptr;
  size_t want = (bytes + 255) & ~size_t(255);
  cudaError_t e = cudaMalloc(&p, want ? want : 256);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    set_error("workspace '%s': cudaMalloc(%zu) failed: %s", name, want, cudaGetErrorString(e));
    return e == cudaErrorMemoryAllocation ? LLIE_E_NOMEM : LLIE_E_CUDA;
  }
  bufs[name] = {p, want};
  ++ws_generation;
  *out = p;
  return LLIE_OK;
}

size_t llie_ctx::workspace_total() const {
  size_t t = 0;
  for (auto& kv : bufs) t += kv.second.second;
  return t;
}

static void msrcr_defaults(llie_msrcr_params* d) {
  memset(d, 0, sizeof(*d));
  d->struct_size = sizeof(*d);
  d->n_scales = 3;
  d->sigma[0] = 15.f; d->sigma[1] = 80.f; d->sigma[2] = 250.f;
  for (int i = 0; i < 3; ++i) d->weight[i] = 1.0f / 3.0f;
  d->truncate = 4.0f;
  d->alpha = 125.f; d->beta = 46.f; d->gain = 0.6f; d->offset = 128.f;
  d->in_scale = 255.f; d->in_bias = 1.f;
  d->blur = LLIE_BLUR_PYRAMID;
  d->finish = LLIE_FINISH_NONE;
  d->finish_k = 2.0f;
}

// Size of the first published layout of llie_msrcr_params (through `blur`); `finish` / `finish_k` came later.
static constexpr size_t kMsrcrParamsV1 = offsetof(llie_msrcr_params, finish);

namespace llie {
// Copies the caller's struct (struct_size bytes: the v1 layout without finish / finish_k, or the current one) into
// a full one; fields the caller's layout does not have keep their defaults.
int msrcr_params_normalise(const llie_msrcr_params* p, llie_msrcr_params* full) {
  LLIE_REQUIRE(p, LLIE_E_ARG, "null params");
  const uint32_t sz = p->struct_size;
  LLIE_REQUIRE(sz >= kMsrcrParamsV1 && sz <= sizeof(*p), LLIE_E_ARG, "llie_msrcr_params.struct_size %u is not in [%zu, %zu]",
               sz, kMsrcrParamsV1, sizeof(*p));
  msrcr_defaults(full);
  memcpy(full, p, sz);
  full->struct_size = sizeof(*full);
  return LLIE_OK;
}
}  // namespace llie

extern "C" {

int llie_version(void) { return LLIE_VERSION; }

const char* llie_last_error(void) { return g_err; }

int llie_device_count(int* count) {
  LLIE_REQUIRE(count, LLIE_E_ARG, "null count");
  int c = 0;
  cudaError_t e = cudaGetDeviceCount(&c);
  if (e != cudaSuccess || c == 0) {
    (void)cudaGetLastError();
    *count = 0;
    set_error("no CUDA device: %s (libllie has no CPU fallback)", e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
    return LLIE_E_CUDA;
  }
  *count = c;
  return LLIE_OK;
}

// The structs are versioned by their leading struct_size: the CALLER states how many bytes its
// struct has, and the library never writes or reads beyond that.
int llie_msrcr_default_params(llie_msrcr_params* p) {
  LLIE_REQUIRE(p, LLIE_E_ARG, "null params");
  const uint32_t sz = p->struct_size;
  LLIE_REQUIRE(sz >= kMsrcrParamsV1 && sz <= sizeof(*p), LLIE_E_ARG,
               "llie_msrcr_params.struct_size %u: set it to sizeof(llie_msrcr_params) (%zu; at least %zu) before the call",
               sz, sizeof(*p), kMsrcrParamsV1);
  llie_msrcr_params d;
  msrcr_defaults(&d);
  d.struct_size = sz;
  memcpy(p, &d, sz);
  return LLIE_OK;
}

int llie_dce_default_params(llie_dce_params* p) {
  LLIE_REQUIRE(p, LLIE_E_ARG, "null params");
  LLIE_REQUIRE(p->struct_size == sizeof(*p), LLIE_E_ARG,
               "llie_dce_params.struct_size %u: set it to sizeof(llie_dce_params) (%zu) before the call", p->struct_size,
               sizeof(*p));
  p->iters = 8;
  p->curve_sign = 1;
  p->precision = LLIE_F32;
  return LLIE_OK;
}

int llie_create(llie_ctx** out, int device) {
  LLIE_REQUIRE(out, LLIE_E_ARG, "null ctx out-pointer");
  *out = nullptr;
  int count = 0;
  LLIE_TRY(llie_device_count(&count));
  LLIE_REQUIRE(device >= 0 && device < count, LLIE_E_ARG, "device %d out of range [0,%d)", device, count);
  DeviceGuard g(device);
  LLIE_REQUIRE(g.ok, LLIE_E_CUDA, "cannot select device %d", device);
  cudaDeviceProp prop;
  LLIE_CUDA_OK(cudaGetDeviceProperties(&prop, device));
  LLIE_REQUIRE(prop.major == 10, LLIE_E_CUDA,
               "device %d is sm_%d%d; libllie is built for sm_100a (B200) only and has no other code path", device,
               prop.major, prop.minor);
  llie_ctx* c = new (std::nothrow) llie_ctx;
  LLIE_REQUIRE(c, LLIE_E_NOMEM, "out of host memory");
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    delete c;
    set_error("cudaStreamCreate failed: %s", cudaGetErrorString(e));
    return LLIE_E_CUDA;
  }
  *out = c;
  return LLIE_OK;
}

int llie_destroy(llie_ctx* ctx) {
  if (!ctx) return LLIE_OK;
  delete ctx->worker;  // joins the ctx's host thread (it is idle: calls are synchronous)
  ctx->worker = nullptr;
  DeviceGuard g(ctx->device);
  cudaDeviceSynchronize();
  sharded_release(ctx);
  for (auto& kv : ctx->bufs) cudaFree(kv.second.first);
  msrcr_plan_free(ctx->plan);
  dce_weights_free(ctx->dce);
  dce_tc_release(ctx);
  if (ctx->aux_stream) { cudaEventDestroy(ctx->aux_ev[0]); cudaEventDestroy(ctx->aux_ev[1]); cudaStreamDestroy(ctx->aux_stream); }
  for (auto& gs : ctx->g_tile)
    if (gs.exec) cudaGraphExecDestroy(gs.exec);
  if (ctx->tile_event) { cudaEventDestroy(ctx->tile_event); cudaEventDestroy(ctx->tile_t0); cudaEventDestroy(ctx->tile_t1); }
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return LLIE_OK;
}

int llie_sync(llie_ctx* ctx) {
  LLIE_REQUIRE(ctx, LLIE_E_ARG, "null ctx");
  DeviceGuard g(ctx->device);
  LLIE_CUDA_OK(cudaStreamSynchronize(ctx->stream));
  return LLIE_OK;
}

int llie_workspace_bytes(llie_ctx* ctx, size_t* bytes) {
  LLIE_REQUIRE(ctx && bytes, LLIE_E_ARG, "null argument");
  *bytes = ctx->workspace_total();
  return LLIE_OK;
}

int llie_launch_count(llie_ctx* ctx, uint64_t* count) {
  LLIE_REQUIRE(ctx && count, LLIE_E_ARG, "null argument");
  *count = ctx->launches;
  return LLIE_OK;
}

int llie_msrcr_tiled_last_ms(llie_ctx* ctx, float* ms) {
  LLIE_REQUIRE(ctx && ms, LLIE_E_ARG, "null argument");
  LLIE_REQUIRE(ctx->tile_ms >= 0.f, LLIE_E_STATE, "no successful llie_msrcr_tiled call has used this ctx");
  *ms = ctx->tile_ms;
  return LLIE_OK;
}

int llie_profile_begin(llie_ctx* ctx) {
  LLIE_REQUIRE(ctx, LLIE_E_ARG, "null ctx");
  DeviceGuard g(ctx->device);
  for (auto& r : ctx->prof) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  ctx->prof.clear();
  ctx->profiling = true;
  return LLIE_OK;
}

int llie_profile_end(llie_ctx* ctx, char* buf, size_t buflen) {
  LLIE_REQUIRE(ctx && buf && buflen > 0, LLIE_E_ARG, "null argument");
  DeviceGuard g(ctx->device);
  ctx->profiling = false;
  LLIE_CUDA_OK(cudaDeviceSynchronize());
  std::map<std::string, std::pair<int, double>> agg;
  for (auto& r : ctx->prof) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) {
      auto& a = agg[r.name];
      a.first++;
      a.second += ms;
    }
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
  }
  ctx->prof.clear();
  std::string s;
  char line[160];
  for (auto& kv : agg) {
    snprintf(line, sizeof(line), "%s:%d:%.6f;", kv.first.c_str(), kv.second.first, kv.second.second);
    s += line;
  }
  LLIE_REQUIRE(s.size() + 1 <= buflen, LLIE_E_ARG, "profile buffer too small (%zu needed)", s.size() + 1);
  memcpy(buf, s.c_str(), s.size() + 1);
  return LLIE_OK;
}

static int check_msrcr_params(const llie_msrcr_params* p) {
  LLIE_REQUIRE(p, LLIE_E_ARG, "null params");
  LLIE_REQUIRE(p->n_scales >= 1 && p->n_scales <= LLIE_MAX_SCALES, LLIE_E_ARG, "n_scales %d out of range", p->n_scales);
  for (int s = 0; s < p->n_scales; ++s)
    LLIE_REQUIRE(p->sigma[s] > 0.f && p->sigma[s] <= 4096.f, LLIE_E_ARG, "sigma[%d]=%g out of range", s, p->sigma[s]);
  LLIE_REQUIRE(p->truncate > 0.f && p->truncate <= 8.f, LLIE_E_ARG, "truncate %g out of range", p->truncate);
  // I = in_scale x + in_bias must be a positive NORMAL fp32 number with room for the products the kernels form
  // (the epilogue takes lg2.approx.ftz, which flushes denormals to zero)
  LLIE_REQUIRE(p->in_bias >= 1e-30f && p->in_bias <= 1e30f && p->in_scale >= 0.f && p->in_scale <= 1e30f, LLIE_E_ARG,
               "in_bias must lie in [1e-30, 1e30] and in_scale in [0, 1e30] (got %g, %g)", p->in_bias, p->in_scale);
  LLIE_REQUIRE(p->blur == LLIE_BLUR_PYRAMID || p->blur == LLIE_BLUR_DIRECT, LLIE_E_ARG, "unknown blur mode %d", p->blur);
  LLIE_REQUIRE(p->finish >= LLIE_FINISH_NONE && p->finish <= LLIE_FINISH_PERCENTILE, LLIE_E_ARG, "unknown finish mode %d", p->finish);
  LLIE_REQUIRE(p->finish == LLIE_FINISH_NONE || p->blur == LLIE_BLUR_PYRAMID, LLIE_E_ARG,
               "global-statistics finishers need the pyramid blur");
  LLIE_REQUIRE(p->finish != LLIE_FINISH_MEANSTD || (p->finish_k > 0.f && p->finish_k <= 16.f), LLIE_E_ARG,
               "finish_k %g out of range (0, 16]", p->finish_k);
  LLIE_REQUIRE(p->finish != LLIE_FINISH_PERCENTILE || (p->finish_k >= 0.f && p->finish_k < 50.f), LLIE_E_ARG,
               "finish_k %g out of range [0, 50) percent", p->finish_k);
  return LLIE_OK;
}

int llie_msrcr_f32(llie_ctx* ctx, const float* in, float* out, int n, int h, int w, int layout,
                   const llie_msrcr_params* params, void* stream) {
  LLIE_REQUIRE(ctx, LLIE_E_ARG, "null ctx");
  LLIE_TRY(check_batch(in, out, n, h, w, layout));
  llie_msrcr_params full;
  LLIE_TRY(msrcr_params_normalise(params, &full));
  params = &full;
  LLIE_TRY(check_msrcr_params(params));
  if (n == 0) return LLIE_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
  int rc = params->blur == LLIE_BLUR_DIRECT ? msrcr_direct_run(ctx, in, out, n, h, w, layout, *params, st)
                                            : msrcr_pyramid_run(ctx, in, out, n, h, w, layout, *params, st);
  LLIE_TRY(rc);
  if (params->finish != LLIE_FINISH_NONE) LLIE_TRY(msrcr_finish_run(ctx, out, n, h, w, layout, *params, st));
  if (!stream) LLIE_CUDA_OK(cudaStreamSynchronize(st));
  return LLIE_OK;
}

static int msrcr_int(llie_ctx* ctx, const void* in, void* out, int iot, int n, int h, int w, int layout,
                     const llie_msrcr_params* params, void* stream) {
  LLIE_REQUIRE(ctx, LLIE_E_ARG, "null ctx");
  LLIE_TRY(check_batch(in, out, n, h, w, layout));
  llie_msrcr_params full;
  LLIE_TRY(msrcr_params_normalise(params, &full));
  params = &full;
  LLIE_TRY(check_msrcr_params(params));
  LLIE_REQUIRE(params->blur == LLIE_BLUR_PYRAMID, LLIE_E_DTYPE, "integer images are supported by the pyramid blur only");
  LLIE_REQUIRE(params->finish == LLIE_FINISH_NONE, LLIE_E_DTYPE, "global-statistics finishers need fp32 images");
  if (n == 0) return LLIE_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
  LLIE_TRY(msrcr_pyramid_run_io(ctx, in, out, iot, n, h, w, layout, *params, st));
  if (!stream) LLIE_CUDA_OK(cudaStreamSynchronize(st));
  return LLIE_OK;
}

int llie_msrcr_u8(llie_ctx* ctx, const uint8_t* in, uint8_t* out, int n, int h, int w, int layout,
                  const llie_msrcr_params* params, void* stream) {
  return msrcr_int(ctx, in, out, kIoU8, n, h, w, layout, params, stream);
}

int llie_msrcr_u16(llie_ctx* ctx, const uint16_t* in, uint16_t* out, int n, int h, int w, int layout,
                   const llie_msrcr_params* params, void* stream) {
  return msrcr_int(ctx, in, out, kIoU16, n, h, w, layout, params, stream);
}

int llie_msrcr_plan_query(float sigma, float truncate, int* stride, double* sigma_low, int* radius_low,
                          double* l1_error) {
  LLIE_REQUIRE(stride && sigma_low && radius_low && l1_error, LLIE_E_ARG, "null out-pointer");
  return msrcr_plan_query(sigma, truncate, stride, sigma_low, radius_low, l1_error);
}

int llie_msrcr_plan_errors(float sigma, float truncate, double* l1_error, double* edge_error, double* line_error) {
  LLIE_REQUIRE(l1_error && edge_error && line_error, LLIE_E_ARG, "null out-pointer");
  return msrcr_plan_detail(sigma, truncate, l1_error, edge_error, line_error, nullptr, 0, nullptr);
}

int llie_msrcr_plan_taps(float sigma, float truncate, float* taps, int capacity, int* count) {
  LLIE_REQUIRE(count, LLIE_E_ARG, "null count");
  LLIE_REQUIRE(capacity >= 0 && (taps || capacity == 0), LLIE_E_ARG, "bad tap buffer");
  return msrcr_plan_detail(sigma, truncate, nullptr, nullptr, nullptr, capacity ? taps : nullptr, capacity, count);
}

int llie_dce_set_weights(llie_ctx* ctx, const float* host_packed, size_t count) {
  LLIE_REQUIRE(ctx, LLIE_E_ARG, "null ctx");
  DeviceGuard g(ctx->device);
  LLIE_CUDA_OK(cudaDeviceSynchronize());  // weights may be in use by queued work
  return dce_weights_upload(ctx, host_packed, count);
}

static int check_dce_params(const llie_dce_params* p) {
  LLIE_REQUIRE(p, LLIE_E_ARG, "null params");
  LLIE_REQUIRE(p->struct_size == sizeof(llie_dce_params), LLIE_E_ARG, "llie_dce_params.struct_size %u != %zu",
               p->struct_size, sizeof(llie_dce_params));
  LLIE_REQUIRE(p->iters >= 0 && p->iters <= 8, LLIE_E_ARG, "iters %d out of range [0,8]", p->iters);
  LLIE_REQUIRE(p->curve_sign == 1 || p->curve_sign == -1, LLIE_E_ARG, "curve_sign must be +1 or -1");
  LLIE_REQUIRE(p->precision == LLIE_F32 || p->precision == LLIE_BF16 || p->precision == LLIE_F32_SIMT, LLIE_E_DTYPE, "unknown precision %d", p->precision);
  return LLIE_OK;
}

int llie_dce_forward(llie_ctx* ctx, const float* in, float* out, float* maps_or_null, int n, int h, int w, int layout,
                     const llie_dce_params* params, void* stream) {
  LLIE_REQUIRE(ctx, LLIE_E_ARG, "null ctx");
  LLIE_TRY(check_batch(in, out, n, h, w, layout));
  LLIE_TRY(check_dce_params(params));
  if (n == 0) return LLIE_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
  int rc = params->precision == LLIE_F32_SIMT ? dce_forward_f32(ctx, in, out, maps_or_null, n, h, w, layout, *params, st)
           : params->precision == LLIE_F32    ? dce_forward_split(ctx, in, out, maps_or_null, n, h, w, layout, *params, st)
                                              : dce_forward_bf16(ctx, in, out, maps_or_null, n, h, w, layout, *params, st);
  LLIE_TRY(rc);
  if (!stream) LLIE_CUDA_OK(cudaStreamSynchronize(st));
  return LLIE_OK;
}

static int dce_int(llie_ctx* ctx, const void* in, void* out, int iot, float* maps_or_null, int n, int h, int w,
                   int layout, const llie_dce_params* params, void* stream) {
  LLIE_REQUIRE(ctx, LLIE_E_ARG, "null ctx");
  LLIE_TRY(check_batch(in, out, n, h, w, layout));
  LLIE_TRY(check_dce_params(params));
  LLIE_REQUIRE(params->precision == LLIE_BF16, LLIE_E_DTYPE, "integer and bfloat16 images are supported by the LLIE_BF16 path only");
  if (n == 0) return LLIE_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
  LLIE_TRY(dce_forward_int(ctx, in, out, iot, maps_or_null, n, h, w, layout, *params, st));
  if (!stream) LLIE_CUDA_OK(cudaStreamSynchronize(st));
  return LLIE_OK;
}

int llie_dce_forward_bf16io(llie_ctx* ctx, const void* in, void* out, float* maps_or_null, int n, int h, int w,
                            int layout, const llie_dce_params* params, void* stream) {
  return dce_int(ctx, in, out, kIoBF16, maps_or_null, n, h, w, layout, params, stream);
}

int llie_dce_forward_u8(llie_ctx* ctx, const uint8_t* in, uint8_t* out, float* maps_or_null, int n, int h, int w,
                        int layout, const llie_dce_params* params, void* stream) {
  return dce_int(ctx, in, out, kIoU8, maps_or_null, n, h, w, layout, params, stream);
}

int llie_dce_forward_u16(llie_ctx* ctx, const uint16_t* in, uint16_t* out, float* maps_or_null, int n, int h, int w,
                         int layout, const llie_dce_params* params, void* stream) {
  return dce_int(ctx, in, out, kIoU16, maps_or_null, n, h, w, layout, params, stream);
}

int llie_debug_dce_activation(llie_ctx* ctx, int which, int frame, float* host_out) {
  LLIE_REQUIRE(ctx, LLIE_E_ARG, "null ctx");
  DeviceGuard g(ctx->device);
  return dce_debug_activation(ctx, which, frame, host_out);
}

int llie_debug_dce_mode(llie_ctx* ctx, int mode) {
  LLIE_REQUIRE(ctx, LLIE_E_ARG, "null ctx");
  LLIE_REQUIRE(mode >= 0 && mode <= 3, LLIE_E_ARG, "mode %d out of range [0,3]", mode);
  return dce_debug_mode(ctx, mode);
}

int llie_dce_curves(llie_ctx* ctx, const float* in, const float* maps, float* out, int n, int h, int w, int layout,
                    const llie_dce_params* params, void* stream) {
  LLIE_REQUIRE(ctx, LLIE_E_ARG, "null ctx");
  LLIE_TRY(check_batch(in, out, n, h, w, layout));
  LLIE_REQUIRE(maps, LLIE_E_ARG, "null maps pointer");
  LLIE_TRY(check_dce_params(params));
  if (n == 0) return LLIE_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
  LLIE_TRY(dce_curves_run(ctx, in, maps, make_view(layout, 24, h, w), out, nullptr, n, h, w, layout, *params, st));
  if (!stream) LLIE_CUDA_OK(cudaStreamSynchronize(st));
  return LLIE_OK;
}

}  // extern "C"
