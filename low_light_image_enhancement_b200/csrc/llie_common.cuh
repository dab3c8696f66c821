// Shared host-side plumbing of libllie.so: error handling, the context object and its
// workspace arena.  SURVEY.md 8(b): the ctx owns streams, workspaces and packed weights;
// every failure becomes a status code, nothing throws across the C boundary.
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>

#include <atomic>
#include <condition_variable>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/llie.h"

namespace llie {

void set_error(const char* fmt, ...);

#define LLIE_CUDA_OK(expr)                                                                      \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      ::llie::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return LLIE_E_CUDA;                                                                       \
    }                                                                                           \
  } while (0)

#define LLIE_TRY(expr)        \
  do {                        \
    int _s = (expr);          \
    if (_s != LLIE_OK) return _s; \
  } while (0)

#define LLIE_REQUIRE(cond, code, ...) \
  do {                                \
    if (!(cond)) {                    \
      ::llie::set_error(__VA_ARGS__); \
      return (code);                  \
    }                                 \
  } while (0)

// Element (c, y, x) of one image lives at base + c*sc + y*sy + x*sx (in elements).
struct ImgView {
  long long sn, sc, sy, sx;
};
inline ImgView make_view(int layout, int c, int h, int w) {
  ImgView v;
  if (layout == LLIE_NHWC) {
    v.sx = c; v.sc = 1; v.sy = (long long)w * c;
  } else {
    v.sx = 1; v.sc = (long long)h * w; v.sy = w;
  }
  v.sn = (long long)h * w * c;
  return v;
}

// One persistent host thread per ctx for the multi-device entry points (llie_enhance_sharded, llie_msrcr_tiled):
// the ~15-25 launches and copies a device needs per call must be issued concurrently with the other devices',
// and spawning a std::thread per call costs more than a whole 8-GPU tiled step.  Jobs run one at a time.
class HostWorker {
 public:
  HostWorker() : th_([this] { loop(); }) {}
  ~HostWorker() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      quit_.store(true, std::memory_order_release);
    }
    cv_.notify_all();
    th_.join();
  }
  void post(std::function<void()> job) {
    {
      std::lock_guard<std::mutex> lk(mu_);
      job_ = std::move(job);
      done_.store(false, std::memory_order_relaxed);
      pending_.store(true, std::memory_order_release);
    }
    cv_.notify_all();
  }
  // Both sides spin briefly before they sleep: a condition-variable wake-up costs 30-60 us, which is most of what a
  // caller of the 8-GPU row-tiled step (0.7 ms of device work) would otherwise wait on top of the devices.
  void wait() {
    for (int i = 0; i < kSpin; ++i) {
      if (done_.load(std::memory_order_acquire)) return;
      relax();
    }
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [this] { return done_.load(std::memory_order_acquire); });
  }

 private:
  static constexpr int kSpin = 6000;  // ~0.2-0.3 ms of pause instructions
  static void relax() {
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#else
    std::this_thread::yield();
#endif
  }
  void loop() {
    for (;;) {
      bool got = false;
      for (int i = 0; i < kSpin && !got; ++i) {
        got = pending_.load(std::memory_order_acquire) || quit_.load(std::memory_order_acquire);
        if (!got) relax();
      }
      std::function<void()> job;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [this] { return pending_.load(std::memory_order_acquire) || quit_.load(std::memory_order_acquire); });
        if (quit_.load(std::memory_order_acquire)) return;
        job = std::move(job_);
        pending_.store(false, std::memory_order_relaxed);
      }
      job();
      {
        std::lock_guard<std::mutex> lk(mu_);
        done_.store(true, std::memory_order_release);
      }
      cv_.notify_all();
    }
  }
  std::mutex mu_;
  std::condition_variable cv_;
  std::function<void()> job_;
  std::atomic<bool> pending_{false}, done_{true}, quit_{false};
  std::thread th_;  // last member: the thread starts after everything above is constructed
};

const char* last_error_text();  // this thread's llie_last_error() text (llie_api.cu)

struct MsrcrPlan;   // msrcr_pyramid.cu
struct DceWeights;  // dce_weights.cu
struct DceTcPack;   // dce_tc.cu: packed weight tiles and activation workspace geometry of the tensor-core path

}  // namespace llie

struct llie_ctx {
  int device = 0;
  int sm_count = 0;
  cudaStream_t stream = nullptr;
  uint64_t launches = 0;
  std::map<std::string, std::pair<void*, size_t>> bufs;  // named workspaces, grown on demand
  uint64_t ws_generation = 0;  // bumped whenever a workspace is (re)allocated: captured graphs hold their addresses
  llie::MsrcrPlan* plan = nullptr;
  llie::DceWeights* dce = nullptr;
  llie::DceTcPack* tc = nullptr;
  // optional per-kernel timing (llie_profile_begin/end): one event pair per launch
  bool profiling = false;
  struct ProfRec { const char* name; cudaEvent_t e0, e1; };
  std::vector<ProfRec> prof;
  cudaEvent_t prof_pending = nullptr;
  cudaEvent_t tile_event = nullptr;  // "my level-C rows are ready" (row-tiled MSRCR)
  cudaEvent_t tile_t0 = nullptr, tile_t1 = nullptr;  // device-side span of this shard in the last tiled call
  float tile_ms = -1.f;
  llie::HostWorker* worker = nullptr;  // created on first multi-device use
  // batch sharder (llie_enhance_sharded): copy streams, per-slot events, pinned stages, last device-side span
  cudaStream_t sh_h2d = nullptr, sh_d2h = nullptr;
  cudaEvent_t sh_ev[3][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};
  cudaEvent_t sh_t0 = nullptr, sh_t1 = nullptr;
  void* sh_pin[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};
  size_t sh_pin_bytes[2][3] = {{0, 0, 0}, {0, 0, 0}};
  float shard_ms = -1.f;
  cudaStream_t aux_stream = nullptr;  // MSRCR: the stride-16 branch runs beside the stride-4 FIRs
  cudaEvent_t aux_ev[2] = {nullptr, nullptr};
  // CUDA graphs of the row-tiled path's launch sequences (phase A; the local tail of phase B), replayed while the
  // key -- pointers, geometry, parameters, workspace generation -- stays the same
  struct GraphSlot { uint64_t key = 0; cudaGraphExec_t exec = nullptr; uint64_t launches = 0; int seen = 0; };
  GraphSlot g_tile[2];

  // Returns a device buffer of at least `bytes` under `name`; reallocates only when growing.
  int workspace(const char* name, size_t bytes, void** out);
  size_t workspace_total() const;
};

namespace llie {

inline void prof_start(llie_ctx* ctx, cudaStream_t st) {
  if (!ctx->profiling) return;
  cudaEventCreate(&ctx->prof_pending);
  cudaEventRecord(ctx->prof_pending, st);
}

inline int check_launch(llie_ctx* ctx, const char* what, cudaStream_t st = nullptr) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("launch of %s failed: %s", what, cudaGetErrorString(e));
    return LLIE_E_CUDA;
  }
  ctx->launches++;
  if (ctx->profiling && ctx->prof_pending) {
    cudaEvent_t e1;
    cudaEventCreate(&e1);
    cudaEventRecord(e1, st);
    ctx->prof.push_back({what, ctx->prof_pending, e1});
    ctx->prof_pending = nullptr;
  }
  return LLIE_OK;
}

// Launch a kernel with error check, launch counting and optional event timing:
//   LLIE_LAUNCH(ctx, stream, "name", kernel<<<grid, block, smem, stream>>>(args...));
#define LLIE_LAUNCH(ctx, st, name, ...)                  \
  do {                                                   \
    ::llie::prof_start(ctx, st);                         \
    __VA_ARGS__;                                         \
    LLIE_TRY(::llie::check_launch(ctx, name, st));       \
  } while (0)

// image sample types of the typed entry points: float in [0, 1], or an integer b meaning b / io_levels
// (kIoBF16: a bfloat16 in [0, 1] -- Zero-DCE with bf16 operands only)
constexpr int kIoF32 = 0, kIoU8 = 1, kIoU16 = 2, kIoBF16 = 3;
inline float io_levels(int iot) { return iot == kIoU8 ? 255.f : (iot == kIoU16 ? 65535.f : 1.f); }
inline size_t io_bytes(int iot) { return iot == kIoU8 ? 1 : ((iot == kIoU16 || iot == kIoBF16) ? 2 : 4); }

// kernels' host entry points (one per .cu)
int msrcr_direct_run(llie_ctx* ctx, const float* in, float* out, int n, int h, int w, int layout,
                     const llie_msrcr_params& p, cudaStream_t st);
int msrcr_finish_run(llie_ctx* ctx, float* out, int n, int h, int w, int layout, const llie_msrcr_params& p,
                     cudaStream_t st);  // msrcr_finish.cu
size_t msrcr_finish_partials_bytes();   // one band / image: 256 partials
size_t msrcr_finish_partial_size();
int msrcr_finish_stats(llie_ctx* ctx, cudaStream_t st, const float* out, int n, int h, int w, int layout, void* d_part);
int msrcr_finish_apply(llie_ctx* ctx, cudaStream_t st, float* out, int n, int h, int w, int layout, const void* d_part,
                       int nparts, double npx, const llie_msrcr_params& p, void* d_li);
size_t msrcr_select_hist_bytes();
size_t msrcr_select_state_bytes();
int msrcr_select_hist(llie_ctx* ctx, cudaStream_t st, const float* band, int h, int w, int pass, const void* d_state, void* d_hist);
int msrcr_select_scan(llie_ctx* ctx, cudaStream_t st, int pass, const void* d_hist_gather, int nsrc, void* d_state, void* d_li,
                      long long npx_total, float k);
int msrcr_stretch(llie_ctx* ctx, cudaStream_t st, float* band, int h, int w, const void* d_li);
int msrcr_pyramid_run_io(llie_ctx* ctx, const void* in, void* out, int iot, int n, int h, int w, int layout,
                         const llie_msrcr_params& p, cudaStream_t st);
int dce_forward_int(llie_ctx* ctx, const void* in, void* out, int iot, float* maps, int n, int h, int w, int layout,
                    const llie_dce_params& p, cudaStream_t st);  // 8- / 16-bit images, bf16 operands
int msrcr_pyramid_run(llie_ctx* ctx, const float* in, float* out, int n, int h, int w, int layout,
                      const llie_msrcr_params& p, cudaStream_t st);
int msrcr_plan_query(float sigma, float truncate, int* stride, double* sigma_low, int* radius_low, double* l1);
int msrcr_plan_detail(float sigma, float truncate, double* l1, double* edge, double* line, float* taps, int capacity,
                      int* count);
void msrcr_plan_free(MsrcrPlan* p);
int msrcr_tiled_run(llie_ctx* const* ctxs, int n_dev, const float* const* in, float* const* out, const int* row0,
                    const int* rows, int H, int W, const llie_msrcr_params& p);

// Runs body(stream) -- a pure sequence of launches / copies -- as a CUDA graph: captured on first use (or when `key`
// changes) and replayed afterwards, so that ~10-20 dependent small launches cost one graph launch on the host and
// run back to back on the device.  Profiling runs (per-kernel events) bypass the graph.
template <class F>
int run_as_graph(llie_ctx* ctx, llie_ctx::GraphSlot& slot, uint64_t key, cudaStream_t st, F&& body) {
  if (ctx->profiling) return body(st);
  if (slot.key != key) {
    // A sequence is captured the SECOND time it is seen: the first run is plain, so that every kernel in it has been
    // loaded (the first launch of a function loads its module lazily, which a capture does not allow) and one-off
    // shapes never pay for an instantiation.
    if (slot.exec) { cudaGraphExecDestroy(slot.exec); slot.exec = nullptr; }
    slot.key = key;
    slot.seen = 1;
    return body(st);
  }
  if (!slot.exec && slot.seen == 1) {
    slot.seen = 2;  // one attempt per key
    const uint64_t l0 = ctx->launches;
    if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
      const int rc = body(st);
      cudaGraph_t graph = nullptr;
      const cudaError_t e = cudaStreamEndCapture(st, &graph);
      slot.launches = ctx->launches - l0;
      ctx->launches = l0;
      if (rc == LLIE_OK && e == cudaSuccess && graph && cudaGraphInstantiate(&slot.exec, graph, 0) != cudaSuccess) slot.exec = nullptr;
      if (graph) cudaGraphDestroy(graph);
    }
    (void)cudaGetLastError();  // a failed capture is not an error of the call: the sequence then runs as plain launches
  }
  if (!slot.exec) return body(st);
  LLIE_CUDA_OK(cudaGraphLaunch(slot.exec, st));
  ctx->launches += slot.launches;
  return LLIE_OK;
}
// The ctx's auxiliary stream (highest priority: its small CTAs take the slots the large launches free) + two events.
inline int ensure_aux_stream(llie_ctx* ctx) {
  if (ctx->aux_stream) return LLIE_OK;
  int prio_lo = 0, prio_hi = 0;
  LLIE_CUDA_OK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
  LLIE_CUDA_OK(cudaStreamCreateWithPriority(&ctx->aux_stream, cudaStreamNonBlocking, prio_hi));
  LLIE_CUDA_OK(cudaEventCreateWithFlags(&ctx->aux_ev[0], cudaEventDisableTiming));
  LLIE_CUDA_OK(cudaEventCreateWithFlags(&ctx->aux_ev[1], cudaEventDisableTiming));
  return LLIE_OK;
}
inline uint64_t fnv1a(const void* data, size_t n, uint64_t h = 1469598103934665603ull) {
  const unsigned char* p = static_cast<const unsigned char*>(data);
  for (size_t i = 0; i < n; ++i) { h ^= p[i]; h *= 1099511628211ull; }
  return h;
}

int msrcr_params_normalise(const llie_msrcr_params* p, llie_msrcr_params* full);  // llie_api.cu
inline HostWorker* ctx_worker(llie_ctx* ctx) {
  if (!ctx->worker) ctx->worker = new HostWorker;
  return ctx->worker;
}
void sharded_release(llie_ctx* ctx);  // llie_sharded.cu
int sharded_ensure_streams(llie_ctx* ctx);  // creates the ctx's copy streams and slot events on first use

int dce_weights_upload(llie_ctx* ctx, const float* host_packed, size_t count);
void dce_weights_free(DceWeights* w);
int dce_forward_f32(llie_ctx* ctx, const float* in, float* out, float* maps, int n, int h, int w, int layout,
                    const llie_dce_params& p, cudaStream_t st);
int dce_forward_bf16(llie_ctx* ctx, const float* in, float* out, float* maps, int n, int h, int w, int layout,
                     const llie_dce_params& p, cudaStream_t st);
int dce_forward_split(llie_ctx* ctx, const float* in, float* out, float* maps, int n, int h, int w, int layout,
                     const llie_dce_params& p, cudaStream_t st);
void dce_tc_release(llie_ctx* ctx);
int dce_debug_activation(llie_ctx* ctx, int which, int frame, float* host_out);
int dce_debug_mode(llie_ctx* ctx, int mode);
int dce_curves_run(llie_ctx* ctx, const float* in, const float* maps, ImgView maps_view, float* out, float* maps_out,
                   int n, int h, int w, int layout, const llie_dce_params& p, cudaStream_t st);

}  // namespace llie
