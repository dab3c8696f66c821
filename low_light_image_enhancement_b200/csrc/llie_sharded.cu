// Batch sharder (SURVEY.md 3.3, 8(b) `llie_enhance_sharded`, 2.2 B9; BASELINE.json#north_star: "shards batches of
// independent images or frames across the 8xB200 box with no collectives").  One C call drives n_dev devices:
// device d processes its own contiguous sub-batch with the unchanged single-device launch sequence on its ctx's
// stream.  Nothing is exchanged between devices; the only synchronisation is one stream sync per device at the end.
// Every ctx has a persistent host thread (HostWorker), so the devices' launches -- and, for HOST buffers, their
// staging copies -- are issued concurrently without returning to Python in between.
// No reference file exists to cite (/root/reference/README.md:1-2).
#include "llie_common.cuh"

#include <cstring>
#include <set>

namespace llie {

namespace {

constexpr int kSlots = 3;                 // staging slots per device of the host pipeline
constexpr size_t kChunkBytes = 48 << 20;  // input bytes per slot (at least one frame)

struct ShardJob {
  int method, iot, n, h, w, layout, where;
  const void* in;
  void* out;
  llie_msrcr_params mp;
  llie_dce_params dp;
};

int run_frames(llie_ctx* ctx, const ShardJob& j, const void* in, void* out, int n, cudaStream_t st) {
  if (j.method == LLIE_METHOD_MSRCR) {
    if (j.iot == LLIE_IO_U8) return llie_msrcr_u8(ctx, (const uint8_t*)in, (uint8_t*)out, n, j.h, j.w, j.layout, &j.mp, st);
    if (j.iot == LLIE_IO_U16) return llie_msrcr_u16(ctx, (const uint16_t*)in, (uint16_t*)out, n, j.h, j.w, j.layout, &j.mp, st);
    return llie_msrcr_f32(ctx, (const float*)in, (float*)out, n, j.h, j.w, j.layout, &j.mp, st);
  }
  if (j.iot == LLIE_IO_U8) return llie_dce_forward_u8(ctx, (const uint8_t*)in, (uint8_t*)out, nullptr, n, j.h, j.w, j.layout, &j.dp, st);
  if (j.iot == LLIE_IO_BF16) return llie_dce_forward_bf16io(ctx, in, out, nullptr, n, j.h, j.w, j.layout, &j.dp, st);
  if (j.iot == LLIE_IO_U16) return llie_dce_forward_u16(ctx, (const uint16_t*)in, (uint16_t*)out, nullptr, n, j.h, j.w, j.layout, &j.dp, st);
  return llie_dce_forward(ctx, (const float*)in, (float*)out, nullptr, n, j.h, j.w, j.layout, &j.dp, st);
}

int ensure_streams(llie_ctx* ctx) {
  if (ctx->sh_t0) return LLIE_OK;
  LLIE_CUDA_OK(cudaStreamCreateWithFlags(&ctx->sh_h2d, cudaStreamNonBlocking));
  LLIE_CUDA_OK(cudaStreamCreateWithFlags(&ctx->sh_d2h, cudaStreamNonBlocking));
  for (int k = 0; k < 3; ++k)
    for (int s = 0; s < kSlots; ++s) LLIE_CUDA_OK(cudaEventCreateWithFlags(&ctx->sh_ev[k][s], cudaEventDisableTiming));
  LLIE_CUDA_OK(cudaEventCreate(&ctx->sh_t1));
  LLIE_CUDA_OK(cudaEventCreate(&ctx->sh_t0));
  return LLIE_OK;
}

bool is_pinned(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    (void)cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}

int pinned_stage(llie_ctx* ctx, int dir, int slot, size_t bytes, void** out) {
  if (ctx->sh_pin_bytes[dir][slot] < bytes) {
    if (ctx->sh_pin[dir][slot]) cudaFreeHost(ctx->sh_pin[dir][slot]);
    ctx->sh_pin[dir][slot] = nullptr;
    ctx->sh_pin_bytes[dir][slot] = 0;
    cudaError_t e = cudaHostAlloc(&ctx->sh_pin[dir][slot], bytes, cudaHostAllocDefault);
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      set_error("cudaHostAlloc(%zu) for a staging buffer failed: %s", bytes, cudaGetErrorString(e));
      return LLIE_E_NOMEM;
    }
    ctx->sh_pin_bytes[dir][slot] = bytes;
  }
  *out = ctx->sh_pin[dir][slot];
  return LLIE_OK;
}

// Device-resident shard: the single-device sequence on the ctx's stream, bracketed by timing events.
int run_device_shard(llie_ctx* ctx, const ShardJob& j) {
  LLIE_TRY(ensure_streams(ctx));
  LLIE_CUDA_OK(cudaEventRecord(ctx->sh_t0, ctx->stream));
  LLIE_TRY(run_frames(ctx, j, j.in, j.out, j.n, ctx->stream));
  LLIE_CUDA_OK(cudaEventRecord(ctx->sh_t1, ctx->stream));
  return LLIE_OK;
}

// Host-resident shard: frames move through kSlots staging slots on three streams -- upload of chunk i, kernels of
// chunk i-1 and download of chunk i-2 overlap; buffer reuse is ordered by stream-side event waits only, so with
// pinned caller memory the host thread never blocks inside the loop.  Pageable memory is staged through pinned
// buffers owned by the ctx (the host memcpy runs on this device's worker thread, in parallel with the others').
int run_host_shard(llie_ctx* ctx, const ShardJob& j) {
  LLIE_TRY(ensure_streams(ctx));
  const size_t frame = (size_t)j.h * j.w * 3 * io_bytes(j.iot);
  const int per = (int)std::max<size_t>(1, std::min<size_t>((size_t)j.n, kChunkBytes / frame));
  const int n_chunks = (j.n + per - 1) / per;
  const bool in_pinned = is_pinned(j.in), out_pinned = is_pinned(j.out);
  void *xd[kSlots], *yd[kSlots], *xp[kSlots] = {nullptr, nullptr, nullptr}, *yp[kSlots] = {nullptr, nullptr, nullptr};
  static const char* kX[kSlots] = {"shard.x0", "shard.x1", "shard.x2"};
  static const char* kY[kSlots] = {"shard.y0", "shard.y1", "shard.y2"};
  const int slots = std::min(kSlots, n_chunks);
  for (int s = 0; s < slots; ++s) {
    LLIE_TRY(ctx->workspace(kX[s], per * frame, &xd[s]));
    LLIE_TRY(ctx->workspace(kY[s], per * frame, &yd[s]));
    if (!in_pinned) LLIE_TRY(pinned_stage(ctx, 0, s, per * frame, &xp[s]));
    if (!out_pinned) LLIE_TRY(pinned_stage(ctx, 1, s, per * frame, &yp[s]));
  }
  cudaEvent_t* ev_in = ctx->sh_ev[0];
  cudaEvent_t* ev_comp = ctx->sh_ev[1];
  cudaEvent_t* ev_out = ctx->sh_ev[2];
  cudaStream_t comp = ctx->stream, h2d = ctx->sh_h2d, d2h = ctx->sh_d2h;
  const char* src_base = static_cast<const char*>(j.in);
  char* dst_base = static_cast<char*>(j.out);
  struct Pending { int slot, lo, hi; };
  std::vector<Pending> pending;
  auto drain = [&](int slot) {  // slot < 0: everything
    size_t k = 0;
    for (size_t i = 0; i < pending.size(); ++i) {
      const Pending& q = pending[i];
      if (slot < 0 || q.slot == slot) memcpy(dst_base + (size_t)q.lo * frame, yp[q.slot], (size_t)(q.hi - q.lo) * frame);
      else pending[k++] = q;
    }
    pending.resize(k);
  };
  LLIE_CUDA_OK(cudaEventRecord(ctx->sh_t0, comp));
  for (int ci = 0; ci < n_chunks; ++ci) {
    const int s = ci % slots, lo = ci * per, hi = std::min(j.n, lo + per);
    const size_t bytes = (size_t)(hi - lo) * frame;
    if (ci >= slots && !out_pinned) {
      LLIE_CUDA_OK(cudaEventSynchronize(ev_out[s]));  // slot s: its previous download is complete
      drain(s);
    }
    const void* src = src_base + (size_t)lo * frame;
    if (!in_pinned) {
      if (ci >= slots) LLIE_CUDA_OK(cudaEventSynchronize(ev_in[s]));  // the previous upload out of xp[s] is complete
      memcpy(xp[s], src, bytes);
      src = xp[s];
    }
    if (ci >= slots) LLIE_CUDA_OK(cudaStreamWaitEvent(h2d, ev_comp[s], 0));  // the kernels that last read xd[s] are done
    LLIE_CUDA_OK(cudaMemcpyAsync(xd[s], src, bytes, cudaMemcpyHostToDevice, h2d));
    LLIE_CUDA_OK(cudaEventRecord(ev_in[s], h2d));
    LLIE_CUDA_OK(cudaStreamWaitEvent(comp, ev_in[s], 0));
    if (ci >= slots) LLIE_CUDA_OK(cudaStreamWaitEvent(comp, ev_out[s], 0));  // the download that last read yd[s] is done
    LLIE_TRY(run_frames(ctx, j, xd[s], yd[s], hi - lo, comp));
    LLIE_CUDA_OK(cudaEventRecord(ev_comp[s], comp));
    LLIE_CUDA_OK(cudaStreamWaitEvent(d2h, ev_comp[s], 0));
    LLIE_CUDA_OK(cudaMemcpyAsync(out_pinned ? (void*)(dst_base + (size_t)lo * frame) : yp[s], yd[s], bytes,
                                 cudaMemcpyDeviceToHost, d2h));
    LLIE_CUDA_OK(cudaEventRecord(ev_out[s], d2h));
    if (!out_pinned) pending.push_back({s, lo, hi});
  }
  LLIE_CUDA_OK(cudaStreamWaitEvent(comp, ev_out[(n_chunks - 1) % slots], 0));  // so that t1 closes the whole pipeline
  LLIE_CUDA_OK(cudaEventRecord(ctx->sh_t1, comp));
  LLIE_CUDA_OK(cudaStreamSynchronize(d2h));
  drain(-1);
  return LLIE_OK;
}

}  // namespace

int sharded_ensure_streams(llie_ctx* ctx) { return ensure_streams(ctx); }

void sharded_release(llie_ctx* ctx) {
  for (int dir = 0; dir < 2; ++dir)
    for (int s = 0; s < kSlots; ++s)
      if (ctx->sh_pin[dir][s]) cudaFreeHost(ctx->sh_pin[dir][s]);
  for (int k = 0; k < 3; ++k)
    for (int s = 0; s < kSlots; ++s)
      if (ctx->sh_ev[k][s]) cudaEventDestroy(ctx->sh_ev[k][s]);
  if (ctx->sh_t0) cudaEventDestroy(ctx->sh_t0);
  if (ctx->sh_t1) cudaEventDestroy(ctx->sh_t1);
  if (ctx->sh_h2d) cudaStreamDestroy(ctx->sh_h2d);
  if (ctx->sh_d2h) cudaStreamDestroy(ctx->sh_d2h);
}

}  // namespace llie

using namespace llie;

extern "C" {

int llie_create_group(llie_ctx** out_ctxs, const int* device_ids, int n_dev) {
  LLIE_REQUIRE(out_ctxs && device_ids, LLIE_E_ARG, "null argument");
  LLIE_REQUIRE(n_dev >= 1 && n_dev <= 64, LLIE_E_ARG, "n_dev %d out of range [1,64]", n_dev);
  for (int d = 0; d < n_dev; ++d) out_ctxs[d] = nullptr;
  for (int d = 0; d < n_dev; ++d) {
    const int rc = llie_create(&out_ctxs[d], device_ids[d]);
    if (rc != LLIE_OK) {  // all or nothing
      for (int e = 0; e < d; ++e) { llie_destroy(out_ctxs[e]); out_ctxs[e] = nullptr; }
      return rc;
    }
  }
  return LLIE_OK;
}

int llie_destroy_group(llie_ctx* const* ctxs, int n_dev) {
  if (!ctxs) return LLIE_OK;
  for (int d = 0; d < n_dev; ++d) llie_destroy(ctxs[d]);
  return LLIE_OK;
}

int llie_enhance_sharded(llie_ctx* const* ctxs, int n_dev, int method, int io_type, const void* const* in_per_dev,
                         void* const* out_per_dev, const int* n_per_dev, int h, int w, int layout, const void* params,
                         int where) {
  LLIE_REQUIRE(ctxs && in_per_dev && out_per_dev && n_per_dev && params, LLIE_E_ARG, "null argument");
  LLIE_REQUIRE(n_dev >= 1 && n_dev <= 64, LLIE_E_ARG, "n_dev %d out of range [1,64]", n_dev);
  LLIE_REQUIRE(method == LLIE_METHOD_MSRCR || method == LLIE_METHOD_ZERO_DCE, LLIE_E_ARG, "unknown method %d", method);
  LLIE_REQUIRE(io_type == LLIE_IO_F32 || io_type == LLIE_IO_U8 || io_type == LLIE_IO_U16 || io_type == LLIE_IO_BF16, LLIE_E_DTYPE,
               "unknown io_type %d", io_type);
  LLIE_REQUIRE(io_type != LLIE_IO_BF16 || method == LLIE_METHOD_ZERO_DCE, LLIE_E_DTYPE, "bfloat16 images are a Zero-DCE format; MSRCR takes float32, uint8 or uint16");
  LLIE_REQUIRE(where == LLIE_SHARD_DEVICE || where == LLIE_SHARD_HOST, LLIE_E_ARG, "unknown buffer location %d", where);
  LLIE_REQUIRE(h > 0 && w > 0 && (long long)h * w <= (1ll << 31), LLIE_E_SHAPE, "bad shape h=%d w=%d", h, w);
  LLIE_REQUIRE(layout == LLIE_NHWC || layout == LLIE_NCHW, LLIE_E_ARG, "unknown layout %d", layout);
  ShardJob proto;
  memset(&proto, 0, sizeof(proto));
  proto.method = method; proto.iot = io_type; proto.h = h; proto.w = w; proto.layout = layout; proto.where = where;
  if (method == LLIE_METHOD_MSRCR) {
    LLIE_TRY(msrcr_params_normalise(static_cast<const llie_msrcr_params*>(params), &proto.mp));
  } else {
    const llie_dce_params* dp = static_cast<const llie_dce_params*>(params);
    LLIE_REQUIRE(dp->struct_size == sizeof(llie_dce_params), LLIE_E_ARG, "llie_dce_params.struct_size %u != %zu",
                 dp->struct_size, sizeof(llie_dce_params));
    proto.dp = *dp;
  }
  std::set<llie_ctx*> uniq;
  std::set<int> devs;
  for (int d = 0; d < n_dev; ++d) {
    LLIE_REQUIRE(ctxs[d], LLIE_E_ARG, "null ctx in device slot %d", d);
    LLIE_REQUIRE(n_per_dev[d] >= 0, LLIE_E_SHAPE, "negative frame count in device slot %d", d);
    LLIE_REQUIRE(n_per_dev[d] == 0 || (in_per_dev[d] && out_per_dev[d]), LLIE_E_ARG, "null image pointer in device slot %d", d);
    LLIE_REQUIRE(n_per_dev[d] == 0 || in_per_dev[d] != out_per_dev[d], LLIE_E_ARG, "in-place operation is not supported (slot %d)", d);
    uniq.insert(ctxs[d]);
    devs.insert(ctxs[d]->device);
  }
  LLIE_REQUIRE((int)uniq.size() == n_dev, LLIE_E_ARG, "every device slot needs its own ctx");
  // distinct devices, or all on one ("virtual devices": the same code path on a one-GPU box, as llie_msrcr_tiled)
  LLIE_REQUIRE(devs.size() == 1 || (int)devs.size() == n_dev, LLIE_E_ARG,
               "ctxs must live on %d distinct devices, or all on one device", n_dev);

  std::vector<int> rc(n_dev, LLIE_OK);
  std::vector<std::string> msg(n_dev);
  std::vector<ShardJob> jobs(n_dev, proto);
  auto body = [&](int d) {
    llie_ctx* ctx = ctxs[d];
    ctx->shard_ms = -1.f;
    if (jobs[d].n == 0) return;
    int prev = -1;
    cudaGetDevice(&prev);
    cudaSetDevice(ctx->device);
    rc[d] = where == LLIE_SHARD_HOST ? run_host_shard(ctx, jobs[d]) : run_device_shard(ctx, jobs[d]);
    if (rc[d] == LLIE_OK) {
      cudaError_t e = cudaStreamSynchronize(ctx->stream);  // the one synchronisation per device
      if (e == cudaSuccess) e = cudaEventElapsedTime(&ctx->shard_ms, ctx->sh_t0, ctx->sh_t1);
      if (e != cudaSuccess) {
        set_error("sharded enhance: device %d: %s", ctx->device, cudaGetErrorString(e));
        rc[d] = LLIE_E_CUDA;
      }
    }
    if (rc[d] != LLIE_OK) {
      msg[d] = last_error_text();
      // A shard that failed half-way may still have copies and kernels in flight that touch the caller's buffers:
      // drain this device's three streams before the caller learns of the error and frees or reuses them (the
      // output of a failed call is undefined).  Errors of the drain itself add nothing to the one being reported.
      for (cudaStream_t s : {ctx->stream, ctx->sh_h2d, ctx->sh_d2h})
        if (s) (void)cudaStreamSynchronize(s);
      (void)cudaGetLastError();
    }
    if (prev >= 0) cudaSetDevice(prev);
  };
  for (int d = 0; d < n_dev; ++d) {
    jobs[d].n = n_per_dev[d];
    jobs[d].in = in_per_dev[d];
    jobs[d].out = out_per_dev[d];
  }
  for (int d = 1; d < n_dev; ++d) ctx_worker(ctxs[d])->post([&body, d] { body(d); });
  body(0);  // the calling thread drives slot 0
  for (int d = 1; d < n_dev; ++d) ctxs[d]->worker->wait();
  for (int d = 0; d < n_dev; ++d)
    if (rc[d] != LLIE_OK) {
      set_error("%s", msg[d].c_str());
      return rc[d];
    }
  return LLIE_OK;
}

int llie_shard_last_ms(llie_ctx* ctx, float* ms) {
  LLIE_REQUIRE(ctx && ms, LLIE_E_ARG, "null argument");
  LLIE_REQUIRE(ctx->shard_ms >= 0.f, LLIE_E_STATE, "no successful llie_enhance_sharded call has used this ctx");
  *ms = ctx->shard_ms;
  return LLIE_OK;
}

}  // extern "C"
