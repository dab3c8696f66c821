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
__global__ void msrcr_stats_fold_kernel(const Partial* __restrict__ part, float2* __restrict__ lo_inv, double npx, int mode, float k) {
  const int img = blockIdx.x, c = threadIdx.x;
  if (c >= 3) return;
  double a = 0, b = 0; float lo = INFINITY, hi = -INFINITY;
  for (int i = 0; i < kStatBlocks; ++i) {
    const Partial& p = part[(long long)img * kStatBlocks + i];
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

}  // namespace

int msrcr_finish_run(llie_ctx* ctx, float* out, int n, int h, int w, int layout, const llie_msrcr_params& p, cudaStream_t st) {
  void *d_part, *d_li;
  LLIE_TRY(ctx->workspace("msrcr.finish.partials", (size_t)n * kStatBlocks * sizeof(Partial), &d_part));
  LLIE_TRY(ctx->workspace("msrcr.finish.lo_inv", (size_t)n * 3 * sizeof(float2), &d_li));
  const ImgView v = make_view(layout, 3, h, w);
  LLIE_LAUNCH(ctx, st, "msrcr_stats_kernel",
              (msrcr_stats_kernel<<<dim3(kStatBlocks, n), kStatThreads, 0, st>>>(out, v, h, w, (Partial*)d_part)));
  LLIE_LAUNCH(ctx, st, "msrcr_stats_fold_kernel",
              (msrcr_stats_fold_kernel<<<n, 32, 0, st>>>((const Partial*)d_part, (float2*)d_li, (double)h * w, p.finish, p.finish_k)));
  const long long npx = (long long)h * w;
  LLIE_LAUNCH(ctx, st, "msrcr_stretch_kernel",
              (msrcr_stretch_kernel<<<dim3((unsigned)((npx + 255) / 256), n), 256, 0, st>>>(out, v, h, w, (const float2*)d_li)));
  return LLIE_OK;
}

}  // namespace llie
