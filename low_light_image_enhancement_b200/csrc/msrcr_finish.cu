// MSRCR global-statistics finishers: per image and per channel, stretch the un-clipped MSRCR
// value v by its min / max or by mean +- k std.  No reference file exists to cite (/root/reference/README.md:1-2);
// the definition is oracle/msrcr.py `finish`.
//
// Three launches on the finished image (which the final kernel wrote un-clipped): per-block partial statistics
// (fixed partition, fp64 block sums: deterministic, no atomics) -> one block per image folds them into (lo, 1/(hi-lo))
// per channel -> in-place stretch.  Algorithmic HBM bytes: 12 (statistics pass) + 24 (stretch) per pixel.
#include "llie_common.cuh"

namespace llie {
namespace {

constexpr int kStatBlocks = 256;   // partial-statistics blocks per image
constexpr int kStatThreads = 256;

struct Partial { double sum[3], sumsq[3]; float mn[3], mx[3]; };

__global__ void __launch_bounds__(kStatThreads)
msrcr_stats_kernel(const float* __restrict__ v, ImgView iv, int H, int W, Partial* __restrict__ part) {
  const long long img = blockIdx.y;
  const long long npx = (long long)H * W;
  const long long per = (npx + kStatBlocks - 1) / kStatBlocks;
  const long long p0 = blockIdx.x * per, p1 = min(p0 + per, npx);
  float s[3] = {0.f, 0.f, 0.f}, q[3] = {0.f, 0.f, 0.f}, mn[3], mx[3];
  double ds[3] = {0, 0, 0}, dq[3] = {0, 0, 0};
#pragma unroll
  for (int c = 0; c < 3; ++c) mn[c] = INFINITY, mx[c] = -INFINITY;
  int cnt = 0;
  for (long long p = p0 + threadIdx.x; p < p1; p += kStatThreads) {
    const int y = (int)(p / W), x = (int)(p - (long long)y * W);
    const float* src = v + img * iv.sn + (long long)y * iv.sy + (long long)x * iv.sx;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float t = __ldg(src + c * iv.sc);
      s[c] += t; q[c] = fmaf(t, t, q[c]);
      mn[c] = fminf(mn[c], t); mx[c] = fmaxf(mx[c], t);
    }
    if (++cnt == 64) {  // fold the fp32 running sums into fp64 before they lose bits
#pragma unroll
      for (int c = 0; c < 3; ++c) { ds[c] += s[c]; dq[c] += q[c]; s[c] = 0.f; q[c] = 0.f; }
      cnt = 0;
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) { ds[c] += s[c]; dq[c] += q[c]; }
  __shared__ double sh_s[3][kStatThreads / 32], sh_q[3][kStatThreads / 32];
  __shared__ float sh_mn[3][kStatThreads / 32], sh_mx[3][kStatThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    for (int o = 16; o > 0; o >>= 1) {
      ds[c] += __shfl_down_sync(0xffffffffu, ds[c], o);
      dq[c] += __shfl_down_sync(0xffffffffu, dq[c], o);
      mn[c] = fminf(mn[c], __shfl_down_sync(0xffffffffu, mn[c], o));
      mx[c] = fmaxf(mx[c], __shfl_down_sync(0xffffffffu, mx[c], o));
    }
    if (lane == 0) { sh_s[c][warp] = ds[c]; sh_q[c][warp] = dq[c]; sh_mn[c][warp] = mn[c]; sh_mx[c][warp] = mx[c]; }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    Partial out;
    for (int c = 0; c < 3; ++c) {
      double a = 0, b = 0; float lo = INFINITY, hi = -INFINITY;
      for (int i = 0; i < kStatThreads / 32; ++i) { a += sh_s[c][i]; b += sh_q[c][i]; lo = fminf(lo, sh_mn[c][i]); hi = fmaxf(hi, sh_mx[c][i]); }
      out.sum[c] = a; out.sumsq[c] = b; out.mn[c] = lo; out.mx[c] = hi;
    }
    part[img * kStatBlocks + blockIdx.x] = out;
  }
}

// lo_inv[img][c] = (lo, 1 / (hi - lo)); a constant channel (hi == lo) maps to 0
__global__ void msrcr_stats_fold_kernel(const Partial* __restrict__ part, int nparts, float2* __restrict__ lo_inv, double npx, int mode, float k) {
  const int img = blockIdx.x, c = threadIdx.x;
  if (c >= 3) return;
  double a = 0, b = 0; float lo = INFINITY, hi = -INFINITY;
  for (int i = 0; i < nparts; ++i) {
    const Partial& p = part[(long long)img * nparts + i];
    a += p.sum[c]; b += p.sumsq[c]; lo = fminf(lo, p.mn[c]); hi = fmaxf(hi, p.mx[c]);
  }
  double l = lo, h = hi;
  if (mode == LLIE_FINISH_MEANSTD) {
    const double mean = a / npx, var = fmax(b / npx - mean * mean, 0.0), sd = sqrt(var);
    l = mean - k * sd; h = mean + k * sd;
  }
  lo_inv[img * 3 + c] = make_float2((float)l, h > l ? (float)(1.0 / (h - l)) : 0.f);
}

__global__ void __launch_bounds__(256)
msrcr_stretch_kernel(float* __restrict__ v, ImgView iv, int H, int W, const float2* __restrict__ lo_inv) {
  const long long img = blockIdx.y;
  const long long p = (long long)blockIdx.x * 256 + threadIdx.x;
  if (p >= (long long)H * W) return;
  const int y = (int)(p / W), x = (int)(p - (long long)y * W);
  float* dst = v + img * iv.sn + (long long)y * iv.sy + (long long)x * iv.sx;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float2 li = __ldg(lo_inv + img * 3 + c);
    dst[c * iv.sc] = __saturatef((dst[c * iv.sc] - li.x) * li.y);
  }
}


// ---- percentile stretch ("simplest colour balance"): exact order statistics by a 3-pass radix select ------------
// Per image and channel, lo = the r-th smallest un-clipped value and hi = the r-th largest, r = floor(N k / 100).
// Keys are the order-preserving 32-bit image of the float; passes take 11, 11 and 10 bits.  Each pass: per-block
// shared-memory histograms of the digit (of the keys that match the prefix selected so far, one histogram per
// target) -> global histogram -> one warp per (channel, target) walks it and extends the prefix.
constexpr int kSelBins = 2048;
struct SelState { unsigned prefix[3][2]; unsigned rank[3][2]; };  // per image: [channel][lo, hi]

__device__ __forceinline__ unsigned sel_key(float f) {
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float sel_value(unsigned k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

template <int PASS>
__global__ void __launch_bounds__(kStatThreads)
msrcr_select_hist_kernel(const float* __restrict__ v, ImgView iv, int H, int W, const SelState* __restrict__ state,
                         unsigned* __restrict__ hist) {
  constexpr int kShift = PASS == 0 ? 21 : (PASS == 1 ? 10 : 0);
  constexpr int kMatchShift = PASS == 1 ? 21 : 10;  // bits above it are already selected
  constexpr unsigned kMask = PASS == 2 ? 1023u : 2047u;
  constexpr int kT = PASS == 0 ? 1 : 2;              // pass 0: both targets share one histogram
  __shared__ unsigned sh[3][kT][kSelBins];
  for (int i = threadIdx.x; i < 3 * kT * kSelBins; i += kStatThreads) (&sh[0][0][0])[i] = 0u;
  const long long img = blockIdx.y;
  unsigned pre[3][2];
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int t = 0; t < 2; ++t) pre[c][t] = PASS == 0 ? 0u : (state[img].prefix[c][t] >> kMatchShift);
  __syncthreads();
  const long long npx = (long long)H * W;
  const long long per = (npx + kStatBlocks - 1) / kStatBlocks;
  const long long p0 = blockIdx.x * per, p1 = min(p0 + per, npx);
  for (long long p = p0 + threadIdx.x; p < p1; p += kStatThreads) {
    const int y = (int)(p / W), x = (int)(p - (long long)y * W);
    const float* src = v + img * iv.sn + (long long)y * iv.sy + (long long)x * iv.sx;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const unsigned k = sel_key(__ldg(src + c * iv.sc));
      const unsigned d = (k >> kShift) & kMask;
      if (PASS == 0) {
        atomicAdd(&sh[c][0][d], 1u);
      } else {
        const unsigned top = k >> kMatchShift;
        if (top == pre[c][0]) atomicAdd(&sh[c][0][d], 1u);
        if (top == pre[c][1]) atomicAdd(&sh[c][kT - 1][d], 1u);
      }
    }
  }
  __syncthreads();
  unsigned* g = hist + img * (3 * 2 * kSelBins);
  for (int i = threadIdx.x; i < 3 * kT * kSelBins; i += kStatThreads) {
    const unsigned cnt = (&sh[0][0][0])[i];
    if (cnt) {
      const int c = i / (kT * kSelBins), t = (i / kSelBins) % kT, b = i % kSelBins;
      atomicAdd(g + (c * 2 + t) * kSelBins + b, cnt);
    }
  }
}

// one warp per (channel, target): find the bin holding the wanted rank, extend the prefix, keep the rank inside it
// hist: nsrc histograms src_stride elements apart (one per device band on the row-tiled path), summed on the fly
template <int PASS>
__global__ void msrcr_select_scan_kernel(const unsigned* __restrict__ hist, int nsrc, long long src_stride,
                                         SelState* __restrict__ state, float2* __restrict__ lo_inv, long long npx, float k) {
  constexpr int kShift = PASS == 0 ? 21 : (PASS == 1 ? 10 : 0);
  const int img = blockIdx.x, lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;  // 6 warps
  const int c = wrp >> 1, t = wrp & 1;
  __shared__ float sh_val[3][2];
  unsigned rank, prefix;
  if (PASS == 0) {
    long long r = (long long)floor((double)npx * (double)k / 100.0);
    r = min(max(r, 0LL), (npx - 1) / 2);
    rank = (unsigned)(t == 0 ? r : npx - 1 - r);
    prefix = 0u;
  } else {
    rank = state[img].rank[c][t]; prefix = state[img].prefix[c][t];
  }
  const unsigned* h0 = hist + (long long)img * (3 * 2 * kSelBins) + (c * 2 + (PASS == 0 ? 0 : t)) * kSelBins;
  auto bin = [&](int b) {
    unsigned n = 0;
    for (int e = 0; e < nsrc; ++e) n += h0[e * src_stride + b];
    return n;
  };
  constexpr int kPer = kSelBins / 32;
  unsigned mine = 0;
  for (int i = 0; i < kPer; ++i) mine += bin(lane * kPer + i);
  unsigned incl = mine;
  for (int o = 1; o < 32; o <<= 1) { const unsigned n = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += n; }
  const unsigned excl = incl - mine;
  const bool has = rank >= excl && rank < incl;   // exactly one lane (the counts sum to more than rank)
  if (has) {
    unsigned acc = excl; int b = lane * kPer;
    for (;; ++b) { const unsigned n = bin(b); if (rank < acc + n) break; acc += n; }
    prefix |= (unsigned)b << kShift; rank -= acc;
    state[img].prefix[c][t] = prefix; state[img].rank[c][t] = rank;
    if (PASS == 2) sh_val[c][t] = sel_value(prefix);
  }
  if (PASS == 2) {
    __syncthreads();
    if (threadIdx.x < 3) {
      const double l = sh_val[threadIdx.x][0], hh = sh_val[threadIdx.x][1];
      lo_inv[img * 3 + threadIdx.x] = make_float2((float)l, hh > l ? (float)(1.0 / (hh - l)) : 0.f);
    }
  }
}

template <int PASS>
int select_pass(llie_ctx* ctx, cudaStream_t st, const float* out, const ImgView& v, int n, int h, int w, SelState* d_state,
                unsigned* d_hist, float2* d_li, float k) {
  LLIE_CUDA_OK(cudaMemsetAsync(d_hist, 0, (size_t)n * 3 * 2 * kSelBins * sizeof(unsigned), st));
  LLIE_LAUNCH(ctx, st, "msrcr_select_hist_kernel",
              (msrcr_select_hist_kernel<PASS><<<dim3(kStatBlocks, n), kStatThreads, 0, st>>>(out, v, h, w, d_state, d_hist)));
  LLIE_LAUNCH(ctx, st, "msrcr_select_scan_kernel",
              (msrcr_select_scan_kernel<PASS><<<n, 192, 0, st>>>(d_hist, 1, 0, d_state, d_li, (long long)h * w, k)));
  return LLIE_OK;
}

}  // namespace

// Percentile finisher across row shards (msrcr_tiled_run): per pass, every device histograms its own band
// (msrcr_select_hist), the histograms are all-gathered by peer copies, and every device runs the same scan over their
// sum (msrcr_select_scan) -- identical prefixes, hence identical (lo, 1/(hi-lo)), everywhere.
size_t msrcr_select_hist_bytes() { return (size_t)3 * 2 * kSelBins * sizeof(unsigned); }
size_t msrcr_select_state_bytes() { return sizeof(SelState); }

int msrcr_select_hist(llie_ctx* ctx, cudaStream_t st, const float* band, int h, int w, int pass, const void* d_state,
                      void* d_hist) {
  const ImgView v = make_view(LLIE_NHWC, 3, h, w);
  LLIE_CUDA_OK(cudaMemsetAsync(d_hist, 0, msrcr_select_hist_bytes(), st));
  const dim3 grid(kStatBlocks, 1);
  if (pass == 0) LLIE_LAUNCH(ctx, st, "msrcr_select_hist_kernel", (msrcr_select_hist_kernel<0><<<grid, kStatThreads, 0, st>>>(band, v, h, w, (const SelState*)d_state, (unsigned*)d_hist)));
  if (pass == 1) LLIE_LAUNCH(ctx, st, "msrcr_select_hist_kernel", (msrcr_select_hist_kernel<1><<<grid, kStatThreads, 0, st>>>(band, v, h, w, (const SelState*)d_state, (unsigned*)d_hist)));
  if (pass == 2) LLIE_LAUNCH(ctx, st, "msrcr_select_hist_kernel", (msrcr_select_hist_kernel<2><<<grid, kStatThreads, 0, st>>>(band, v, h, w, (const SelState*)d_state, (unsigned*)d_hist)));
  return LLIE_OK;
}

int msrcr_select_scan(llie_ctx* ctx, cudaStream_t st, int pass, const void* d_hist_gather, int nsrc, void* d_state,
                      void* d_li, long long npx_total, float k) {
  const long long stride = (long long)(msrcr_select_hist_bytes() / sizeof(unsigned));
  if (pass == 0) LLIE_LAUNCH(ctx, st, "msrcr_select_scan_kernel", (msrcr_select_scan_kernel<0><<<1, 192, 0, st>>>((const unsigned*)d_hist_gather, nsrc, stride, (SelState*)d_state, (float2*)d_li, npx_total, k)));
  if (pass == 1) LLIE_LAUNCH(ctx, st, "msrcr_select_scan_kernel", (msrcr_select_scan_kernel<1><<<1, 192, 0, st>>>((const unsigned*)d_hist_gather, nsrc, stride, (SelState*)d_state, (float2*)d_li, npx_total, k)));
  if (pass == 2) LLIE_LAUNCH(ctx, st, "msrcr_select_scan_kernel", (msrcr_select_scan_kernel<2><<<1, 192, 0, st>>>((const unsigned*)d_hist_gather, nsrc, stride, (SelState*)d_state, (float2*)d_li, npx_total, k)));
  return LLIE_OK;
}

int msrcr_stretch(llie_ctx* ctx, cudaStream_t st, float* band, int h, int w, const void* d_li) {
  const ImgView v = make_view(LLIE_NHWC, 3, h, w);
  const long long npx = (long long)h * w;
  LLIE_LAUNCH(ctx, st, "msrcr_stretch_kernel",
              (msrcr_stretch_kernel<<<dim3((unsigned)((npx + 255) / 256), 1), 256, 0, st>>>(band, v, h, w, (const float2*)d_li)));
  return LLIE_OK;
}

// The two halves of the min/max and mean/std finishers, also used by the row-tiled path (msrcr_tiled_run), where the
// partials of every device's band are gathered before the fold.
size_t msrcr_finish_partials_bytes() { return kStatBlocks * sizeof(Partial); }
size_t msrcr_finish_partial_size() { return sizeof(Partial); }

int msrcr_finish_stats(llie_ctx* ctx, cudaStream_t st, const float* out, int n, int h, int w, int layout, void* d_part) {
  const ImgView v = make_view(layout, 3, h, w);
  LLIE_LAUNCH(ctx, st, "msrcr_stats_kernel",
              (msrcr_stats_kernel<<<dim3(kStatBlocks, n), kStatThreads, 0, st>>>(out, v, h, w, (Partial*)d_part)));
  return LLIE_OK;
}

// nparts partials per image; npx = pixels per image the partials cover (the whole image, not the band)
int msrcr_finish_apply(llie_ctx* ctx, cudaStream_t st, float* out, int n, int h, int w, int layout, const void* d_part,
                       int nparts, double npx, const llie_msrcr_params& p, void* d_li) {
  const ImgView v = make_view(layout, 3, h, w);
  LLIE_LAUNCH(ctx, st, "msrcr_stats_fold_kernel",
              (msrcr_stats_fold_kernel<<<n, 32, 0, st>>>((const Partial*)d_part, nparts, (float2*)d_li, npx, p.finish, p.finish_k)));
  const long long band_px = (long long)h * w;
  LLIE_LAUNCH(ctx, st, "msrcr_stretch_kernel",
              (msrcr_stretch_kernel<<<dim3((unsigned)((band_px + 255) / 256), n), 256, 0, st>>>(out, v, h, w, (const float2*)d_li)));
  return LLIE_OK;
}

int msrcr_finish_run(llie_ctx* ctx, float* out, int n, int h, int w, int layout, const llie_msrcr_params& p, cudaStream_t st) {
  void *d_part, *d_li;
  LLIE_TRY(ctx->workspace("msrcr.finish.partials", (size_t)n * kStatBlocks * sizeof(Partial), &d_part));
  LLIE_TRY(ctx->workspace("msrcr.finish.lo_inv", (size_t)n * 3 * sizeof(float2), &d_li));
  const ImgView v = make_view(layout, 3, h, w);
  if (p.finish == LLIE_FINISH_PERCENTILE) {
    void *d_state, *d_hist;
    LLIE_TRY(ctx->workspace("msrcr.finish.select_state", (size_t)n * sizeof(SelState), &d_state));
    LLIE_TRY(ctx->workspace("msrcr.finish.select_hist", (size_t)n * 3 * 2 * kSelBins * sizeof(unsigned), &d_hist));
    LLIE_TRY(select_pass<0>(ctx, st, out, v, n, h, w, (SelState*)d_state, (unsigned*)d_hist, (float2*)d_li, p.finish_k));
    LLIE_TRY(select_pass<1>(ctx, st, out, v, n, h, w, (SelState*)d_state, (unsigned*)d_hist, (float2*)d_li, p.finish_k));
    LLIE_TRY(select_pass<2>(ctx, st, out, v, n, h, w, (SelState*)d_state, (unsigned*)d_hist, (float2*)d_li, p.finish_k));
    const long long npx = (long long)h * w;
    LLIE_LAUNCH(ctx, st, "msrcr_stretch_kernel",
                (msrcr_stretch_kernel<<<dim3((unsigned)((npx + 255) / 256), n), 256, 0, st>>>(out, v, h, w, (const float2*)d_li)));
  } else {
    LLIE_TRY(msrcr_finish_stats(ctx, st, out, n, h, w, layout, d_part));
    LLIE_TRY(msrcr_finish_apply(ctx, st, out, n, h, w, layout, d_part, kStatBlocks, (double)h * w, p, d_li));
  }
  return LLIE_OK;
}

}  // namespace llie
