// MSRCR, fast path (LLIE_BLUR_PYRAMID): every Gaussian surround is evaluated as a cascade
//     prefilter+decimate  ->  small Gaussian at low resolution  ->  Gaussian-weighted upsample
// whose cost per pixel does not depend on sigma, and the upsample is fused with the per-pixel
// log / ratio / colour-restoration / gain-offset epilogue (north_star kernel (1); SURVEY.md 8(a)
// units a1-a4; spec SURVEY.md 9.1).  No reference file exists to cite (/root/reference/README.md:1-2).
//
// Why a cascade: a direct 4-sigma FIR costs 33 kFLOP/pixel and tops out at 3.7 % of the HBM
// roofline.  Gaussians compose (variances add), a Gaussian-prefiltered image
// can be decimated with aliasing below 1e-9, and a unit-sigma Gaussian sampled on a unit grid is
// a partition of unity to 5e-9, so it doubles as the interpolator.  Two levels are kept:
//   level A: stride 2   (prefilter sigma 2, radius 8, on I = in_scale*x + in_bias)
//   level B: stride 16  (level A prefiltered with sigma 8, radius 32, decimated by 8)
// A scale with 4 <= sigma < 40 is blurred on A, a scale with sigma >= 40 on B.  The low-res sigma
// and radius are chosen by a host-side planner that minimises the L1 distance between the
// cascade's effective kernel and the oracle's truncated Gaussian (tools/pyramid_tune.py is the
// NumPy twin of that planner).  Edge clamping is exact: every level stores enough samples beyond
// the image border to reach the region where the clamped, prefiltered signal is constant.
//
// HBM traffic: the image is read twice (12 B/px each) and written once (12 B/px); level A adds
// ~3 B/px per touch, level B is negligible.  SURVEY.md 8(d)'s two-pass figure is 108 B/px.
#include "fir.cuh"

#include <cmath>
#include <mutex>

namespace llie {

namespace {

constexpr int kT = 4;         // upsample half-width in low-res samples (2*kT+1 taps)
constexpr int kRA = 8;        // level-A prefilter radius (full-res pixels), sigma 2
constexpr int kPA = 4;        // level-A samples stored before pixel 0
constexpr int kRB = 32;       // level-B prefilter radius (level-A samples), sigma 8
constexpr int kPB = 5;        // level-B samples stored before pixel 0
constexpr double kSigA = 2.0, kSigB = 8.0, kSigUp = 1.0;

inline int level_a_size(int n0) { return (n0 + 8) / 2 + 9; }
inline int level_b_size(int n0) { return (n0 + 86) / 16 + 10; }

std::vector<double> gauss_taps(double sigma, int r, double offs = 0.0) {
  std::vector<double> k(2 * r + 1);
  double sum = 0;
  for (int t = -r; t <= r; ++t) {
    double d = t - offs;
    sum += (k[t + r] = std::exp(-d * d / (2 * sigma * sigma)));
  }
  for (double& v : k) v /= sum;
  return k;
}

// weights of the Gaussian upsampler for phase ph of stride s: taps j0-kT .. j0+kT
std::vector<double> upsample_weights(int s, int ph) {
  double shift = (ph >= s / 2 ? 1.0 : 0.0) - (double)ph / s;  // (j0 - P) - X/s
  std::vector<double> w(2 * kT + 1);
  double sum = 0;
  for (int t = 0; t <= 2 * kT; ++t) {
    double d = shift - kT + t;
    sum += (w[t] = std::exp(-d * d / (2 * kSigUp * kSigUp)));
  }
  for (double& v : w) v /= sum;
  return w;
}

// ---- planner ------------------------------------------------------------------------------
struct ScalePlan {
  int level;          // 0 = A (stride 2), 1 = B (stride 16)
  int stride;
  double sigma_low;   // in low-res samples
  int radius_low;
  double l1;          // L1 distance of the effective kernel to the oracle kernel
  std::vector<float> taps;
};

// full-res impulse response of "prefilter + decimate" seen from one low-res sample
std::vector<double> prefilter_fullres(int level) {
  std::vector<double> ga = gauss_taps(kSigA, kRA);
  if (level == 0) return ga;
  std::vector<double> gb = gauss_taps(kSigB, kRB);
  std::vector<double> g(2 * (2 * kRB + kRA) + 1, 0.0);  // support +-(2*32+8)
  const int c = 2 * kRB + kRA;
  for (int m = -kRB; m <= kRB; ++m)
    for (int t = -kRA; t <= kRA; ++t) g[c + 2 * m + t] += gb[m + kRB] * ga[t + kRA];
  return g;
}

double cascade_l1(int s, const std::vector<double>& gpre, const std::vector<std::vector<double>>& wup,
                  const std::vector<double>& klow, const std::vector<double>& kor) {
  const int rl = ((int)klow.size() - 1) / 2, rp = ((int)gpre.size() - 1) / 2, ro = ((int)kor.size() - 1) / 2;
  const int nc = 2 * (rl + kT) + 1;              // composite low-res kernel w * klow, centred on j0
  const int span = s * (rl + kT + 1) + rp + s;   // |delta| bound of the effective kernel
  std::vector<double> c(nc), e(2 * span + 1);
  double total = 0;
  for (int ph = 0; ph < s; ++ph) {
    std::fill(c.begin(), c.end(), 0.0);
    for (int t = 0; t <= 2 * kT; ++t)
      for (int m = 0; m <= 2 * rl; ++m) c[t + m] += wup[ph][t] * klow[m];
    std::fill(e.begin(), e.end(), 0.0);
    // low-res sample j0 + q sits at full-res offset s*(q + [ph >= s/2]) - ph from the pixel
    const int base = s * (ph >= s / 2 ? 1 : 0) - ph;
    for (int q = -(rl + kT); q <= rl + kT; ++q) {
      const double cq = c[q + rl + kT];
      const int off = base + s * q + span;
      for (int t = -rp; t <= rp; ++t) e[off + t] += cq * gpre[t + rp];
    }
    double l1 = 0;
    for (int d = -span; d <= span; ++d) {
      double ko = (d >= -ro && d <= ro) ? kor[d + ro] : 0.0;
      l1 += std::fabs(e[d + span] - ko);
    }
    total += l1;
  }
  return total / s;
}

int plan_scale(double sigma, double truncate, ScalePlan* out) {
  const int level = sigma >= 40.0 ? 1 : 0;
  const int s = level ? 16 : 2;
  const double var_pre = level ? (kSigA * kSigA + 4.0 * kSigB * kSigB) : kSigA * kSigA;
  const double var_up = kSigUp * kSigUp * s * s;
  const double rest = sigma * sigma - var_pre - var_up;
  if (rest < s * s * 1.0) {
    set_error("pyramid blur needs sigma >= 4 (got %g); use blur=direct for small scales", sigma);
    return LLIE_E_ARG;
  }
  const double sl0 = std::sqrt(rest) / s;
  const int ro = (int)std::ceil(truncate * sigma);
  const std::vector<double> kor = gauss_taps(sigma, ro);
  const std::vector<double> gpre = prefilter_fullres(level);
  std::vector<std::vector<double>> wup(s);
  for (int ph = 0; ph < s; ++ph) wup[ph] = upsample_weights(s, ph);

  double best_l1 = 1e30, best_sl = sl0;
  int best_rl = 1;
  const int rl_mid = ro / s;
  for (int rl = std::max(1, rl_mid - 2); rl <= rl_mid + 1; ++rl) {
    // golden-section search of sigma_low in [0.97, 1.03] * sl0
    double a = 0.97 * sl0, b = 1.03 * sl0;
    const double gr = 0.6180339887498949;
    double x1 = b - gr * (b - a), x2 = a + gr * (b - a);
    double f1 = cascade_l1(s, gpre, wup, gauss_taps(x1, rl), kor);
    double f2 = cascade_l1(s, gpre, wup, gauss_taps(x2, rl), kor);
    for (int it = 0; it < 36; ++it) {
      if (f1 < f2) {
        b = x2; x2 = x1; f2 = f1;
        x1 = b - gr * (b - a);
        f1 = cascade_l1(s, gpre, wup, gauss_taps(x1, rl), kor);
      } else {
        a = x1; x1 = x2; f1 = f2;
        x2 = a + gr * (b - a);
        f2 = cascade_l1(s, gpre, wup, gauss_taps(x2, rl), kor);
      }
    }
    const double x = f1 < f2 ? x1 : x2, f = f1 < f2 ? f1 : f2;
    if (f < best_l1) { best_l1 = f; best_sl = x; best_rl = rl; }
  }
  out->level = level;
  out->stride = s;
  out->sigma_low = best_sl;
  out->radius_low = best_rl;
  out->l1 = best_l1;
  std::vector<double> k = gauss_taps(best_sl, best_rl);
  out->taps.assign(k.begin(), k.end());
  return LLIE_OK;
}

// plans are pure functions of (sigma, truncate); memoise them process-wide
std::mutex g_plan_mu;
std::map<std::pair<float, float>, ScalePlan> g_plan_cache;

int get_scale_plan(float sigma, float truncate, ScalePlan* out) {
  std::lock_guard<std::mutex> lk(g_plan_mu);
  auto key = std::make_pair(sigma, truncate);
  auto it = g_plan_cache.find(key);
  if (it == g_plan_cache.end()) {
    ScalePlan sp;
    LLIE_TRY(plan_scale(sigma, truncate, &sp));
    it = g_plan_cache.emplace(key, std::move(sp)).first;
  }
  *out = it->second;
  return LLIE_OK;
}

// ---- kernels ------------------------------------------------------------------------------
struct TapsA { float g[2 * kRA + 1]; };

// Level A: A[c][jy][jx] = sum_{ty,tx} g[ty] g[tx] I[clamp(2(jy-P)+ty-8)][clamp(2(jx-P)+tx-8)][c]
// A CTA produces a 32x16 tile of A from an 80x48 pixel patch staged in shared memory
// (row filter at the decimated columns, then column filter at the decimated rows).
// in: image rows [in_row0, in_row0+in_rows) of an H-row image; out: A rows [a_row0, a_row0+a_rows).
constexpr int kDAW = 32, kDAH = 16, kDPW = 2 * kDAW + 2 * kRA, kDPH = 2 * kDAH + 2 * kRA;  // 80 x 48

__global__ void __launch_bounds__(256)
down2_kernel(const float* __restrict__ x, ImgView v, int H, int W, int in_row0, float* __restrict__ A, int hA_buf,
             int wA, int a_row0, float in_scale, float in_bias, const __grid_constant__ TapsA taps) {
  extern __shared__ __align__(16) float s_dyn[];
  float (*s_in)[kDPH][kDPW] = reinterpret_cast<float (*)[kDPH][kDPW]>(s_dyn);
  float (*s_tmp)[kDPH][kDAW] = reinterpret_cast<float (*)[kDPH][kDAW]>(s_dyn + 3 * kDPH * kDPW);
  const int tid = threadIdx.x;
  const long long img = blockIdx.z;
  const int jx0 = blockIdx.x * kDAW, jyb0 = blockIdx.y * kDAH;  // jyb0: buffer row of A
  const int X0 = 2 * (jx0 - kPA) - kRA, Y0 = 2 * (a_row0 + jyb0 - kPA) - kRA;
  const float* src = x + img * v.sn;
  if (v.sx == 3) {  // NHWC: walk (row, x*3+c) so that consecutive threads read consecutive floats
    for (int i = tid; i < kDPH * kDPW * 3; i += 256) {
      int row = i / (kDPW * 3), e = i - row * (kDPW * 3);
      int px = e / 3, c = e - px * 3;
      int Y = min(max(Y0 + row, 0), H - 1) - in_row0, X = min(max(X0 + px, 0), W - 1);
      s_in[c][row][px] = fmaf(in_scale, __ldg(src + (long long)Y * v.sy + (long long)X * 3 + c), in_bias);
    }
  } else {
    for (int i = tid; i < kDPH * kDPW * 3; i += 256) {
      int c = i / (kDPH * kDPW), rem = i - c * (kDPH * kDPW);
      int row = rem / kDPW, px = rem - row * kDPW;
      int Y = min(max(Y0 + row, 0), H - 1) - in_row0, X = min(max(X0 + px, 0), W - 1);
      s_in[c][row][px] = fmaf(in_scale, __ldg(src + c * v.sc + (long long)Y * v.sy + (long long)X * v.sx), in_bias);
    }
  }
  __syncthreads();
  for (int i = tid; i < 3 * kDPH * kDAW; i += 256) {
    int jx = i % kDAW, row = (i / kDAW) % kDPH, c = i / (kDAW * kDPH);
    const float* p = &s_in[c][row][2 * jx];
    float acc = 0.f;
#pragma unroll
    for (int t = 0; t <= 2 * kRA; ++t) acc = fmaf(taps.g[t], p[t], acc);
    s_tmp[c][row][jx] = acc;
  }
  __syncthreads();
  for (int i = tid; i < 3 * kDAH * kDAW; i += 256) {
    int jx = i % kDAW, jy = (i / kDAW) % kDAH, c = i / (kDAW * kDAH);
    if (jx0 + jx >= wA || jyb0 + jy >= hA_buf) continue;
    float acc = 0.f;
#pragma unroll
    for (int t = 0; t <= 2 * kRA; ++t) acc = fmaf(taps.g[t], s_tmp[c][2 * jy + t][jx], acc);
    A[((img * 3 + c) * hA_buf + jyb0 + jy) * (long long)wA + jx0 + jx] = acc;
  }
}

// Level B, row pass: T[plane][row][jx] = sum_t g[t] A[plane][row][clamp(8(jx-PB)+PA+t-32)]
__global__ void down8_rows_kernel(const float* __restrict__ A, float* __restrict__ T, const float* __restrict__ g,
                                  long long rows_total, int wA, int wB) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= rows_total * wB) return;
  long long row = i / wB;
  int jx = (int)(i - row * wB);
  const float* src = A + row * wA;
  const int c0 = 8 * (jx - kPB) + kPA - kRB;
  float acc = 0.f;
  for (int t = 0; t <= 2 * kRB; ++t) acc = fmaf(g[t], __ldg(src + min(max(c0 + t, 0), wA - 1)), acc);
  T[i] = acc;
}

// Level B, column pass: B[plane][jy][jx] = sum_t g[t] T[plane][clamp(8(jy-PB)+PA+t-32) - a_row0][jx]
__global__ void down8_cols_kernel(const float* __restrict__ T, float* __restrict__ B, const float* __restrict__ g,
                                  int planes, int hA, int hA_buf, int a_row0, int wB, int hB_buf, int b_row0) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= (long long)planes * hB_buf * wB) return;
  int jx = (int)(i % wB);
  int jyb = (int)((i / wB) % hB_buf);
  long long plane = i / ((long long)wB * hB_buf);
  const float* src = T + plane * (long long)hA_buf * wB + jx;
  const int r0 = 8 * (b_row0 + jyb - kPB) + kPA - kRB;
  float acc = 0.f;
  for (int t = 0; t <= 2 * kRB; ++t)
    acc = fmaf(g[t], __ldg(src + (long long)(min(max(r0 + t, 0), hA - 1) - a_row0) * wB), acc);
  B[i] = acc;
}

// Final pass: Gaussian upsample of every scale's low-res surround + MSRCR epilogue.
struct FinalParams {
  int n_scales;
  int level[LLIE_MAX_SCALES];
  float weight[LLIE_MAX_SCALES];
  const float* map[LLIE_MAX_SCALES];  // planar [n][3][rows_buf][wL]
  int rows_buf[2], wL[2], row0[2];    // per level: buffer rows, width, level row held by buffer row 0
  float log_alpha, beta, gain, offset, in_scale, in_bias;
  float wup_a[2][2 * kT + 1];
  float wup_b[16][2 * kT + 1];
};

constexpr int kFX = 64, kFY = 32;                      // output tile
constexpr int kSrcRowsA = kFY / 2 + 2 * kT + 1;        // 25
constexpr int kSrcColsA = kFX / 2 + 2 * kT + 1;        // 41
constexpr int kSrcRowsB = kFY / 16 + 2 * kT + 1;       // 11
constexpr int kSrcColsB = kFX / 16 + 2 * kT + 1;       // 13

__global__ void __launch_bounds__(256)
msrcr_final_kernel(const float* __restrict__ x, float* __restrict__ out, ImgView v, int rows, int W, int row_g0,
                   const __grid_constant__ FinalParams p) {
  __shared__ float s_src[3 * kSrcRowsA * kSrcColsA];   // low-res patch of the current scale
  __shared__ float s_h[3 * kSrcRowsA * kFX];           // horizontally upsampled rows
  const int tid = threadIdx.x;
  const int lx = tid & (kFX - 1), yg = tid >> 6;       // thread: column lx, rows yg*8 .. yg*8+7
  const int x0 = blockIdx.x * kFX, yb0 = blockIdx.y * kFY;  // yb0: row inside this buffer
  const int Y0 = row_g0 + yb0;                         // image row of the tile (multiple of 32)
  const long long img = blockIdx.z;
  const int X = x0 + lx;
  const bool x_ok = X < W;

  float lI[8][3], acc[8][3];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const int yb = yb0 + yg * 8 + q;
    const bool ok = x_ok && yb < rows;
    const long long base = img * v.sn + (long long)yb * v.sy + (long long)X * v.sx;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      lI[q][c] = ok ? fmaf(p.in_scale, __ldg(x + base + c * v.sc), p.in_bias) : 1.f;  // holds I for now
      acc[q][c] = 0.f;
    }
  }

  for (int k = 0; k < p.n_scales; ++k) {
    const int lv = p.level[k];
    const int s = lv ? 16 : 2, P = lv ? kPB : kPA;
    const int sr = lv ? kSrcRowsB : kSrcRowsA, sc = lv ? kSrcColsB : kSrcColsA;
    const int wL = p.wL[lv], hbuf = p.rows_buf[lv];
    const int col0 = x0 / s + P - kT, rowb0 = Y0 / s + P - kT - p.row0[lv];
    const float* m = p.map[k] + img * 3ll * hbuf * wL;
    __syncthreads();  // previous scale's readers are done with s_src / s_h
    for (int i = tid; i < 3 * sr * sc; i += 256) {
      int cc = i % sc, rr = (i / sc) % sr, c = i / (sc * sr);
      int gr = min(rowb0 + rr, hbuf - 1), gc = min(col0 + cc, wL - 1);
      s_src[i] = __ldg(m + ((long long)c * hbuf + gr) * wL + gc);
    }
    __syncthreads();
    {  // stage 1: horizontal upsample; this thread's column lx for every (channel, low-res row)
      const int ph = X % s;
      const float* wx = lv ? p.wup_b[ph] : p.wup_a[ph];
      float w[2 * kT + 1];
#pragma unroll
      for (int t = 0; t <= 2 * kT; ++t) w[t] = wx[t];
      const int lc = (lx + s / 2) / s;  // local column of tap 0
      for (int cr = yg; cr < 3 * sr; cr += 4) {
        const float* src = s_src + cr * sc + lc;
        float a = 0.f;
#pragma unroll
        for (int t = 0; t <= 2 * kT; ++t) a = fmaf(w[t], src[t], a);
        s_h[cr * kFX + lx] = a;
      }
    }
    __syncthreads();
    // stage 2: vertical upsample of this thread's 8 rows, then accumulate -w_k ln S
    const float wk = p.weight[k];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int yl = yg * 8 + q;
      const int ph = (Y0 + yl) % s;
      const float* wy = lv ? p.wup_b[ph] : p.wup_a[ph];
      const int lr = (yl + s / 2) / s;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float* src = s_h + (c * sr + lr) * kFX + lx;
        float S = 0.f;
#pragma unroll
        for (int t = 0; t <= 2 * kT; ++t) S = fmaf(wy[t], src[t * kFX], S);
        acc[q][c] = fmaf(-wk, __logf(S), acc[q][c]);
      }
    }
  }

  float wsum = 0.f;
  for (int k = 0; k < p.n_scales; ++k) wsum += p.weight[k];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const int yb = yb0 + yg * 8 + q;
    if (!x_ok || yb >= rows) continue;
    const long long base = img * v.sn + (long long)yb * v.sy + (long long)X * v.sx;
    const float lsum = __logf(lI[q][0] + lI[q][1] + lI[q][2]);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float l = __logf(lI[q][c]);
      const float R = fmaf(wsum, l, acc[q][c]);
      const float Cc = p.beta * (p.log_alpha + l - lsum);
      const float o = fmaf(p.gain * Cc, R, p.offset) * (1.0f / 255.0f);
      out[base + c * v.sc] = fminf(fmaxf(o, 0.f), 1.f);
    }
  }
}

}  // namespace

// ---- plan object held by the ctx ------------------------------------------------------------
struct MsrcrPlan {
  llie_msrcr_params key;
  std::vector<ScalePlan> scales;
  float* d_taps = nullptr;  // [gB(65) | scale taps ...]
  std::vector<int> tap_off;
  TapsA ga;
  FinalParams fp;
};

void msrcr_plan_free(MsrcrPlan* p) {
  if (!p) return;
  cudaFree(p->d_taps);
  delete p;
}

int msrcr_plan_query(float sigma, float truncate, int* stride, double* sigma_low, int* radius_low, double* l1) {
  ScalePlan sp;
  LLIE_TRY(get_scale_plan(sigma, truncate, &sp));
  *stride = sp.stride;
  *sigma_low = sp.sigma_low;
  *radius_low = sp.radius_low;
  *l1 = sp.l1;
  return LLIE_OK;
}

static int ensure_plan(llie_ctx* ctx, const llie_msrcr_params& p, cudaStream_t st) {
  if (ctx->plan && memcmp(&ctx->plan->key, &p, sizeof(p)) == 0) return LLIE_OK;
  MsrcrPlan* pl = new MsrcrPlan;
  pl->key = p;
  std::vector<float> host;
  {
    std::vector<double> gb = gauss_taps(kSigB, kRB);
    for (double v : gb) host.push_back((float)v);
    host.resize((host.size() + 3) & ~size_t(3));
  }
  for (int k = 0; k < p.n_scales; ++k) {
    ScalePlan sp;
    int rc = get_scale_plan(p.sigma[k], p.truncate, &sp);
    if (rc != LLIE_OK) { delete pl; return rc; }
    pl->tap_off.push_back((int)host.size());
    host.insert(host.end(), sp.taps.begin(), sp.taps.end());
    host.resize((host.size() + 3) & ~size_t(3));
    pl->scales.push_back(std::move(sp));
  }
  {
    std::vector<double> ga = gauss_taps(kSigA, kRA);
    for (int t = 0; t <= 2 * kRA; ++t) pl->ga.g[t] = (float)ga[t];
  }
  FinalParams& fp = pl->fp;
  memset(&fp, 0, sizeof(fp));
  fp.n_scales = p.n_scales;
  for (int k = 0; k < p.n_scales; ++k) { fp.level[k] = pl->scales[k].level; fp.weight[k] = p.weight[k]; }
  fp.log_alpha = (float)std::log((double)p.alpha);
  fp.beta = p.beta; fp.gain = p.gain; fp.offset = p.offset; fp.in_scale = p.in_scale; fp.in_bias = p.in_bias;
  for (int ph = 0; ph < 2; ++ph) {
    auto w = upsample_weights(2, ph);
    for (int t = 0; t <= 2 * kT; ++t) fp.wup_a[ph][t] = (float)w[t];
  }
  for (int ph = 0; ph < 16; ++ph) {
    auto w = upsample_weights(16, ph);
    for (int t = 0; t <= 2 * kT; ++t) fp.wup_b[ph][t] = (float)w[t];
  }
  cudaError_t e = cudaMalloc(&pl->d_taps, host.size() * sizeof(float));
  if (e == cudaSuccess) e = cudaMemcpy(pl->d_taps, host.data(), host.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    set_error("msrcr plan upload failed: %s", cudaGetErrorString(e));
    msrcr_plan_free(pl);
    return LLIE_E_CUDA;
  }
  if (ctx->plan) {
    cudaStreamSynchronize(st);
    msrcr_plan_free(ctx->plan);
  }
  ctx->plan = pl;
  return LLIE_OK;
}

int msrcr_pyramid_run(llie_ctx* ctx, const float* in, float* out, int n, int h, int w, int layout,
                      const llie_msrcr_params& p, cudaStream_t st) {
  LLIE_TRY(ensure_plan(ctx, p, st));
  MsrcrPlan* pl = ctx->plan;
  const int hA = level_a_size(h), wA = level_a_size(w), hB = level_b_size(h), wB = level_b_size(w);
  bool any_a = false, any_b = false;
  int n_a = 0, n_b = 0;
  for (auto& s : pl->scales) (s.level ? (any_b = true, ++n_b) : (any_a = true, ++n_a));
  const size_t a_elems = (size_t)n * 3 * hA * wA, b_elems = (size_t)n * 3 * hB * wB;
  void *dA, *dAt, *dAblur, *dBt, *dB, *dBt2, *dBblur;
  LLIE_TRY(ctx->workspace("msrcr.A", a_elems * 4, &dA));
  LLIE_TRY(ctx->workspace("msrcr.A.tmp", a_elems * 4, &dAt));
  LLIE_TRY(ctx->workspace("msrcr.A.blur", a_elems * 4 * (n_a ? n_a : 1), &dAblur));
  LLIE_TRY(ctx->workspace("msrcr.B.rows", (size_t)n * 3 * hA * wB * 4, &dBt));
  LLIE_TRY(ctx->workspace("msrcr.B", b_elems * 4, &dB));
  LLIE_TRY(ctx->workspace("msrcr.B.tmp", b_elems * 4, &dBt2));
  LLIE_TRY(ctx->workspace("msrcr.B.blur", b_elems * 4 * (n_b ? n_b : 1), &dBblur));

  const ImgView v = make_view(layout, 3, h, w);
  const int planes = n * 3;
  {
    dim3 grid((wA + kDAW - 1) / kDAW, (hA + kDAH - 1) / kDAH, n);
    constexpr size_t sm = (size_t)3 * kDPH * (kDPW + kDAW) * sizeof(float);
    LLIE_CUDA_OK(cudaFuncSetAttribute(down2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    LLIE_LAUNCH(ctx, st, "down2_kernel", down2_kernel<<<grid, 256, sm, st>>>(in, v, h, w, 0, (float*)dA, hA, wA, 0, p.in_scale, p.in_bias, pl->ga));
  }
  if (any_b) {
    const float* gB = pl->d_taps;
    long long rows_total = (long long)planes * hA;
    LLIE_LAUNCH(ctx, st, "down8_rows_kernel", down8_rows_kernel<<<(unsigned)((rows_total * wB + 255) / 256), 256, 0, st>>>((const float*)dA, (float*)dBt, gB,
                                                                                  rows_total, wA, wB));
    long long nb = (long long)planes * hB * wB;
    LLIE_LAUNCH(ctx, st, "down8_cols_kernel", down8_cols_kernel<<<(unsigned)((nb + 255) / 256), 256, 0, st>>>((const float*)dBt, (float*)dB, gB, planes, hA, hA, 0,
                                                                     wB, hB, 0));
  }
  FinalParams fp = pl->fp;
  fp.rows_buf[0] = hA; fp.wL[0] = wA; fp.row0[0] = 0;
  fp.rows_buf[1] = hB; fp.wL[1] = wB; fp.row0[1] = 0;
  int ia = 0, ib = 0;
  for (int k = 0; k < p.n_scales; ++k) {
    const ScalePlan& sp = pl->scales[k];
    const float* taps = pl->d_taps + pl->tap_off[k];
    const int hL = sp.level ? hB : hA, wL = sp.level ? wB : wA;
    const float* src = sp.level ? (const float*)dB : (const float*)dA;
    float* tmp = sp.level ? (float*)dBt2 : (float*)dAt;
    float* dst = sp.level ? (float*)dBblur + (size_t)(ib++) * b_elems : (float*)dAblur + (size_t)(ia++) * a_elems;
    const int r = sp.radius_low;
    size_t sm = (size_t)(128 + 2 * r + 3) * sizeof(float);
    dim3 gx((unsigned)((long long)planes * hL), (wL + 127) / 128);
    LLIE_LAUNCH(ctx, st, "fir_x_kernel", fir_x_kernel<false><<<gx, 128, sm, st>>>(src, tmp, taps, r, wL, 1, 1.f, 0.f));
    dim3 gy((wL + 127) / 128, (hL + 3) / 4, planes);
    LLIE_LAUNCH(ctx, st, "fir_y_kernel", fir_y_kernel<<<gy, 128, 0, st>>>(tmp, dst, taps, r, hL, wL, 0, hL, 0, hL));
    fp.map[k] = dst;
  }
  {
    dim3 grid((w + kFX - 1) / kFX, (h + kFY - 1) / kFY, n);
    LLIE_LAUNCH(ctx, st, "msrcr_final_kernel", msrcr_final_kernel<<<grid, 256, 0, st>>>(in, out, v, h, w, 0, fp));
  }
  (void)any_a;
  return LLIE_OK;
}

}  // namespace llie
