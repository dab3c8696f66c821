// MSRCR, fast path (LLIE_BLUR_PYRAMID): host-side planner and launch sequence of the decimated
// Gaussian cascade in msrcr_kernels.cuh (north_star kernel (1); SURVEY.md 8(a) units a1-a4; spec
// SURVEY.md 9.1).  No reference file exists to cite (/root/reference/README.md:1-2).
//
// Why a cascade: a direct 4-sigma FIR costs 33 kFLOP/pixel and tops out at 3.7 % of the HBM
// roofline.  Gaussians compose (variances add) and a Gaussian-prefiltered image
// can be decimated with aliasing below 1e-8, so every surround is evaluated at stride 4 or 16 and
// brought back by interpolation.  The planner picks, per scale, the low-resolution Gaussian
// (sigma_low, radius_low) that minimises the L1 distance between the cascade's effective kernel
// and the oracle's truncated Gaussian.
//
// HBM traffic per pixel: the image is read twice (12 B each) and written once (12 B); the stride-4
// intermediates add ~3 B (V) + ~3 B (level-C maps).  SURVEY.md 8(d)'s two-pass figure is 108 B/px.
#include "fir.cuh"
#include "msrcr_kernels.cuh"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>

namespace llie {

namespace {

std::vector<double> gauss_taps(double sigma, int r) {
  std::vector<double> k(2 * r + 1);
  double sum = 0;
  for (int t = -r; t <= r; ++t) sum += (k[t + r] = std::exp(-(double)t * t / (2 * sigma * sigma)));
  for (double& v : k) v /= sum;
  return k;
}

// weights of the B -> C Gaussian upsampler for phase ph: taps j0-4 .. j0+4, j0 - u = [ph>=2] - ph/4
void up_weights(int ph, double (&w)[2 * kTU + 1]) {
  const double shift = (ph >= 2 ? 1.0 : 0.0) - ph / 4.0;
  double sum = 0;
  for (int t = 0; t <= 2 * kTU; ++t) {
    const double d = shift - kTU + t;
    sum += (w[t] = std::exp(-d * d / 2.0));
  }
  for (double& v : w) v /= sum;
}

void lagrange_weights(int T, double frac, double* w) {  // offsets -(T/2-1) .. T/2
  for (int i = 0; i < T; ++i) {
    const double oi = i - (T / 2 - 1);
    double v = 1;
    for (int m = 0; m < T; ++m) {
      if (m == i) continue;
      const double om = m - (T / 2 - 1);
      v *= (frac - om) / (oi - om);
    }
    w[i] = v;
  }
}

// ---- planner ------------------------------------------------------------------------------
// A finitely supported 1-D signal on an integer grid: v[i] sits at index lo + i.
struct Sig {
  int lo = 0;
  std::vector<double> v;
};
Sig conv(const Sig& a, const std::vector<double>& k) {  // k centred (radius (len-1)/2)
  const int r = ((int)k.size() - 1) / 2;
  Sig o;
  o.lo = a.lo - r;
  o.v.assign(a.v.size() + 2 * r, 0.0);
  for (size_t i = 0; i < a.v.size(); ++i)
    for (int t = 0; t <= 2 * r; ++t) o.v[i + t] += a.v[i] * k[t];
  return o;
}
// adjoint of "filter with g, decimate by 4": weights on coarse samples -> weights on fine samples
Sig spread4(const Sig& a, const std::vector<double>& g) {
  const int r = ((int)g.size() - 1) / 2;
  Sig o;
  o.lo = 4 * a.lo - r;
  o.v.assign(4 * (a.v.size() - 1) + 2 * r + 1, 0.0);
  for (size_t i = 0; i < a.v.size(); ++i)
    for (int t = 0; t <= 2 * r; ++t) o.v[4 * i + t] += a.v[i] * g[t];
  return o;
}

struct ScalePlan {
  int level;         // 0 = C (stride 4), 1 = B (stride 16)
  int stride;
  double sigma_low;  // in low-res samples
  int radius_low;
  double l1;         // L1 distance of the effective kernel to the oracle kernel (phase average)
  double edge, line; // relative surround errors of a black pixel beside a white half-plane / 1-px line (KernelErr)
  std::vector<float> taps;
};

// weights on level-C samples (index j at position 4j) that the final Lagrange step applies for pixel ph
Sig lagrange_sig(int T, int ph) {
  Sig s;
  const int jc0 = ph >> 2, f = ph & 3;
  if (f == 0) {
    s.lo = jc0;
    s.v = {1.0};
    return s;
  }
  s.lo = jc0 - (T / 2 - 1);
  s.v.resize(T);
  lagrange_weights(T, f / 4.0, s.v.data());
  return s;
}
// weights on level-B samples (index j at position 16j) equivalent to weights on level-C samples
Sig through_upsampler(const Sig& c) {
  Sig b;
  int lo = 1 << 30, hi = -(1 << 30);
  for (size_t i = 0; i < c.v.size(); ++i) {
    const int jc = c.lo + (int)i;
    const int fl = (int)std::floor(jc / 4.0), ph = jc - 4 * fl;
    const int j0 = fl + (ph >= 2 ? 1 : 0);
    lo = std::min(lo, j0 - kTU);
    hi = std::max(hi, j0 + kTU);
  }
  b.lo = lo;
  b.v.assign(hi - lo + 1, 0.0);
  for (size_t i = 0; i < c.v.size(); ++i) {
    const int jc = c.lo + (int)i;
    const int fl = (int)std::floor(jc / 4.0), ph = jc - 4 * fl;
    const int j0 = fl + (ph >= 2 ? 1 : 0);
    double w[2 * kTU + 1];
    up_weights(ph, w);
    for (int t = 0; t <= 2 * kTU; ++t) b.v[j0 - kTU + t - lo] += c.v[i] * w[t];
  }
  return b;
}

double cascade_l1(int level, const std::vector<Sig>& left, const std::vector<double>& gd, const std::vector<double>& klow,
                  const std::vector<double>& kor) {
  const int ro = ((int)kor.size() - 1) / 2;
  double total = 0;
  for (size_t ph = 0; ph < left.size(); ++ph) {
    Sig e = conv(left[ph], klow);
    if (level) e = spread4(e, gd);
    e = spread4(e, gd);  // now on pixels; the pixel itself is index ph
    double l1 = 0;
    const int lo = std::min(e.lo, (int)ph - ro), hi = std::max(e.lo + (int)e.v.size() - 1, (int)ph + ro);
    for (int xx = lo; xx <= hi; ++xx) {
      const double ev = (xx >= e.lo && xx < e.lo + (int)e.v.size()) ? e.v[xx - e.lo] : 0.0;
      const int d = xx - (int)ph;
      const double ko = (d >= -ro && d <= ro) ? kor[d + ro] : 0.0;
      l1 += std::fabs(ev - ko);
    }
    total += l1;
  }
  return total / left.size();
}

constexpr double kMinSigma = 12.0;  // below this the stride-4 grid is too coarse for 1e-4 parity
constexpr double kLevelBSigma = 40.0;
// Contrast the planner designs for: I = in_scale x + in_bias with the default 255 / 1, i.e. a pure black pixel
// (I = 1) next to pure white (I = 256).
constexpr double kDesignContrast = 255.0;

// Error figures of an effective kernel against the oracle's (all phases; both sides):
//   l1   phase-averaged sum |dk|                                    (fine texture, noise)
//   edge max over distance d of rho |sum_{t >= d} dk| / (1 + rho tail_o(d))  = relative error of the surround S
//        of a black pixel at distance d from a white half-plane (rho = 255: I = 1 next to I = 256)
//   line max of rho |dk(d)| / (1 + rho k_o(d))                      = the same for a 1-pixel white line
struct KernelErr { double l1 = 0, edge = 0, line = 0; };

// Unit responses: h[ph] = weights on PIXELS (index = absolute position, the pixel itself is `ph`) of a unit
// low-resolution tap at low-res offset 0, for the pixel of phase ph.  A low-res tap at offset j shifts it by s j.
std::vector<Sig> unit_responses(int level) {
  const int s = level ? 16 : 4;
  const std::vector<double> gd = gauss_taps(4.0, kRD);
  std::vector<Sig> h;
  for (int ph = 0; ph < s; ++ph) {
    Sig c = lagrange_sig(level ? 4 : 6, ph);
    Sig e = level ? through_upsampler(c) : c;
    if (level) e = spread4(e, gd);
    h.push_back(spread4(e, gd));
  }
  return h;
}

// dense effective kernel of phase ph on offsets -R..R (relative to the pixel) for symmetric low-res taps x[0..rl]
void effective_dense(const Sig& h, int ph, int s, const std::vector<double>& x, int R, std::vector<double>* out) {
  out->assign(2 * R + 1, 0.0);
  const int rl = (int)x.size() - 1;
  for (int j = -rl; j <= rl; ++j) {
    const double xv = x[j < 0 ? -j : j];
    if (xv == 0.0) continue;
    // absolute position a = h.lo + i + s j  ->  offset t = a - ph
    for (size_t i = 0; i < h.v.size(); ++i) {
      const int t = h.lo + (int)i + s * j - ph;
      if (t >= -R && t <= R) (*out)[t + R] += xv * h.v[i];
    }
  }
}

KernelErr kernel_errors(int level, const std::vector<Sig>& h, const std::vector<double>& x, const std::vector<double>& kor) {
  const int s = level ? 16 : 4, rl = (int)x.size() - 1, ro = ((int)kor.size() - 1) / 2;
  int ext = 0;
  for (auto& q : h) ext = std::max(ext, std::max(std::abs(q.lo), std::abs(q.lo + (int)q.v.size())));
  const int R = std::max(ro, s * rl + ext) + s;
  const int L = 2 * R + 1;
  std::vector<double> ko(L, 0.0), tail_r(L + 1, 0.0), tail_l(L + 1, 0.0), e;
  for (int t = -ro; t <= ro; ++t) ko[t + R] = kor[t + ro];
  for (int i = L - 1; i >= 0; --i) tail_r[i] = tail_r[i + 1] + ko[i];
  for (int i = 0; i < L; ++i) tail_l[i + 1] = tail_l[i] + ko[i];
  KernelErr r;
  const double rho = kDesignContrast;
  for (int ph = 0; ph < s; ++ph) {
    effective_dense(h[ph], ph, s, x, R, &e);
    double l1 = 0, acc = 0;
    for (int i = L - 1; i >= 0; --i) {  // white half-plane on the right, starting at offset i - R
      const double dk = e[i] - ko[i];
      l1 += std::fabs(dk);
      acc += dk;
      r.edge = std::max(r.edge, rho * std::fabs(acc) / (1.0 + rho * tail_r[i]));
      r.line = std::max(r.line, rho * std::fabs(dk) / (1.0 + rho * ko[i]));
    }
    acc = 0;
    for (int i = 0; i < L; ++i) {  // ... on the left
      acc += e[i] - ko[i];
      r.edge = std::max(r.edge, rho * std::fabs(acc) / (1.0 + rho * tail_l[i + 1]));
    }
    r.l1 += l1 / s;
  }
  return r;
}

// v1 family: a truncated Gaussian (sigma_low, r_low), sigma_low by golden-section search on the L1 distance.
void plan_taps_gauss(int level, double sigma, int ro, const std::vector<Sig>& left, const std::vector<double>& gd,
                     const std::vector<double>& kor, std::vector<double>* taps) {
  const int s = level ? 16 : 4;
  const double var_fixed = level ? (16.0 + 256.0 + 256.0) : 16.0;  // prefilters (+ upsampler), pixel^2
  const double sl0 = std::sqrt(sigma * sigma - var_fixed) / s;
  double best_l1 = 1e30, best_sl = sl0;
  int best_rl = 1;
  const int rl_mid = ro / s;
  for (int rl = std::max(1, rl_mid - 2); rl <= rl_mid + 1; ++rl) {
    double a = 0.97 * sl0, b = 1.03 * sl0;  // golden-section search of sigma_low
    const double gr = 0.6180339887498949;
    double x1 = b - gr * (b - a), x2 = a + gr * (b - a);
    double f1 = cascade_l1(level, left, gd, gauss_taps(x1, rl), kor);
    double f2 = cascade_l1(level, left, gd, gauss_taps(x2, rl), kor);
    for (int it = 0; it < 32; ++it) {
      if (f1 < f2) {
        b = x2; x2 = x1; f2 = f1;
        x1 = b - gr * (b - a);
        f1 = cascade_l1(level, left, gd, gauss_taps(x1, rl), kor);
      } else {
        a = x1; x1 = x2; f1 = f2;
        x2 = a + gr * (b - a);
        f2 = cascade_l1(level, left, gd, gauss_taps(x2, rl), kor);
      }
    }
    const double xb = f1 < f2 ? x1 : x2, fb = f1 < f2 ? f1 : f2;
    if (fb < best_l1) { best_l1 = fb; best_sl = xb; best_rl = rl; }
  }
  *taps = gauss_taps(best_sl, best_rl);
}

// v2: free-form symmetric low-res taps x[0..rl] by weighted linear least squares.  The effective kernel is linear
// in the taps; the objective is the squared error of (a) the one-sided cumulative sums of the kernel (the response
// to a straight edge), each weighted by rho / (1 + rho tail) -- the relative error it causes in the surround of a
// black pixel beside a white half-plane -- and (b) the kernel itself weighted by rho / (1 + rho k_o) (a 1-pixel
// line), subject to sum(taps) = 1.  The v1 family only minimised the L1 distance, which left the oracle's hard
// 4-sigma truncation smeared over ~23 px at stride 16: 5.8e-4 in the output on a black-to-white edge against
// 1.3e-4 (sigma = 15), 1.0e-4 (80), 3.7e-5 (250) now.  What remains is the floor of a stride-4 grid: a jump of
// k_o(r) per pixel cannot be placed more finely than the grid.
bool plan_taps_ls(int level, int rl, const std::vector<Sig>& h, const std::vector<double>& kor, std::vector<double>* x_out) {
  const int s = level ? 16 : 4, ro = ((int)kor.size() - 1) / 2, n = rl + 1;
  int ext = 0;
  for (auto& q : h) ext = std::max(ext, std::max(std::abs(q.lo), std::abs(q.lo + (int)q.v.size())));
  const int R = std::max(ro, s * rl + ext) + s;
  const int L = 2 * R + 1;
  const double rho = kDesignContrast;
  std::vector<double> ko(L, 0.0), tail_r(L + 1, 0.0), tail_l(L + 1, 0.0);
  for (int t = -ro; t <= ro; ++t) ko[t + R] = kor[t + ro];
  for (int i = L - 1; i >= 0; --i) tail_r[i] = tail_r[i + 1] + ko[i];
  for (int i = 0; i < L; ++i) tail_l[i + 1] = tail_l[i] + ko[i];
  std::vector<double> M((size_t)n * n, 0.0), v(n, 0.0);
  std::vector<double> A((size_t)L * n), row(n), unit(n, 0.0), col;
  for (int ph = 0; ph < s; ++ph) {
    // A[i][j] = d e(offset i - R) / d x_j
    for (int j = 0; j < n; ++j) {
      unit.assign(n, 0.0);
      unit[j] = 1.0;
      effective_dense(h[ph], ph, s, unit, R, &col);
      for (int i = 0; i < L; ++i) A[(size_t)i * n + j] = col[i];
    }
    auto add_row = [&](const double* a, double b, double w) {
      if (w == 0.0) return;
      for (int j = 0; j < n; ++j) {
        const double aw = a[j] * w * w;
        if (aw == 0.0) continue;
        v[j] += aw * b;
        double* m = &M[(size_t)j * n];
        for (int k = j; k < n; ++k) m[k] += aw * a[k];
      }
    };
    row.assign(n, 0.0);
    for (int i = L - 1; i >= 0; --i) {  // cumulative from the right
      for (int j = 0; j < n; ++j) row[j] += A[(size_t)i * n + j];
      add_row(row.data(), tail_r[i], rho / (1.0 + rho * tail_r[i]));
      add_row(&A[(size_t)i * n], ko[i], rho / (1.0 + rho * ko[i]));
    }
    row.assign(n, 0.0);
    for (int i = 0; i < L; ++i) {  // cumulative from the left
      for (int j = 0; j < n; ++j) row[j] += A[(size_t)i * n + j];
      add_row(row.data(), tail_l[i + 1], rho / (1.0 + rho * tail_l[i + 1]));
    }
  }
  for (int j = 0; j < n; ++j)
    for (int k = 0; k < j; ++k) M[(size_t)j * n + k] = M[(size_t)k * n + j];
  double tr = 0;
  for (int j = 0; j < n; ++j) tr += M[(size_t)j * n + j];
  for (int j = 0; j < n; ++j) M[(size_t)j * n + j] += 1e-9 * tr / n;  // Tikhonov: the far taps are weakly determined
  // KKT system [M c; c^T 0] [x; mu] = [v; 1], c = (1, 2, 2, ...): Gaussian elimination with partial pivoting
  const int m = n + 1;
  std::vector<long double> K((size_t)m * (m + 1), 0.0L);
  for (int j = 0; j < n; ++j) {
    for (int k = 0; k < n; ++k) K[(size_t)j * (m + 1) + k] = M[(size_t)j * n + k];
    const long double c = j == 0 ? 1.0L : 2.0L;
    K[(size_t)j * (m + 1) + n] = c;
    K[(size_t)n * (m + 1) + j] = c;
    K[(size_t)j * (m + 1) + m] = v[j];
  }
  K[(size_t)n * (m + 1) + m] = 1.0L;
  for (int c = 0; c < m; ++c) {
    int piv = c;
    for (int r2 = c + 1; r2 < m; ++r2)
      if (fabsl(K[(size_t)r2 * (m + 1) + c]) > fabsl(K[(size_t)piv * (m + 1) + c])) piv = r2;
    if (fabsl(K[(size_t)piv * (m + 1) + c]) < 1e-300L) return false;
    if (piv != c)
      for (int k = 0; k <= m; ++k) std::swap(K[(size_t)piv * (m + 1) + k], K[(size_t)c * (m + 1) + k]);
    for (int r2 = c + 1; r2 < m; ++r2) {
      const long double f = K[(size_t)r2 * (m + 1) + c] / K[(size_t)c * (m + 1) + c];
      if (f == 0.0L) continue;
      for (int k = c; k <= m; ++k) K[(size_t)r2 * (m + 1) + k] -= f * K[(size_t)c * (m + 1) + k];
    }
  }
  std::vector<long double> sol(m);
  for (int r2 = m - 1; r2 >= 0; --r2) {
    long double a = K[(size_t)r2 * (m + 1) + m];
    for (int k = r2 + 1; k < m; ++k) a -= K[(size_t)r2 * (m + 1) + k] * sol[k];
    sol[r2] = a / K[(size_t)r2 * (m + 1) + r2];
  }
  x_out->resize(n);
  for (int j = 0; j < n; ++j) {
    (*x_out)[j] = (double)sol[j];
    if (!std::isfinite((*x_out)[j])) return false;
  }
  return true;
}

constexpr int kMaxLsRadius = 96;  // the least-squares fit is O(L n^2) per phase; beyond this (sigma > ~1400) keep v1

int plan_scale(double sigma, double truncate, ScalePlan* out) {
  if (sigma < kMinSigma) {
    set_error("pyramid blur needs sigma >= %g (got %g); use blur=direct for small scales", kMinSigma, sigma);
    return LLIE_E_ARG;
  }
  const int level = sigma >= kLevelBSigma ? 1 : 0;
  const int s = level ? 16 : 4;
  const int ro = (int)std::ceil(truncate * sigma);
  const std::vector<double> kor = gauss_taps(sigma, ro);
  const std::vector<double> gd = gauss_taps(4.0, kRD);
  std::vector<Sig> left;
  for (int ph = 0; ph < s; ++ph) {
    Sig c = lagrange_sig(level ? 4 : 6, ph);
    left.push_back(level ? through_upsampler(c) : c);
  }
  const std::vector<Sig> h = unit_responses(level);
  auto half = [](const std::vector<double>& k) {  // centred taps -> x[0..rl]
    const int rl = ((int)k.size() - 1) / 2;
    return std::vector<double>(k.begin() + rl, k.end());
  };
  // candidate 1: the Gaussian family (v1)
  std::vector<double> kg;
  plan_taps_gauss(level, sigma, ro, left, gd, kor, &kg);
  std::vector<double> best = half(kg);
  KernelErr best_err = kernel_errors(level, h, best, kor);
  // candidate 2: weighted least squares with a few more taps than the oracle's radius covers
  // a few more taps than the oracle's radius covers: stride 16 costs nothing (1/256 of the pixels); at stride 4
  // one extra tap keeps the FIR at the same number of 8-tap steps as the radius itself (2 r + 8 rounded up to 8)
  const int rl_ls = (ro + s - 1) / s + (level ? 5 : 1);
  if (rl_ls <= kMaxLsRadius) {
    std::vector<double> x;
    if (plan_taps_ls(level, rl_ls, h, kor, &x)) {
      const KernelErr e = kernel_errors(level, h, x, kor);
      if (std::max(e.edge, e.line) < std::max(best_err.edge, best_err.line) && e.l1 < 4.0 * best_err.l1 + 1e-4) {
        best = x;
        best_err = e;
      }
    }
  }
  const int rl = (int)best.size() - 1;
  std::vector<double> k(2 * rl + 1);
  double m2 = 0, m0 = 0;
  for (int t = -rl; t <= rl; ++t) {
    k[t + rl] = best[t < 0 ? -t : t];
    m0 += k[t + rl];
    m2 += k[t + rl] * t * t;
  }
  out->level = level;
  out->stride = s;
  out->sigma_low = std::sqrt(std::max(m2 / m0, 0.0));  // second moment of the taps (they are free-form now)
  out->radius_low = rl;
  out->l1 = best_err.l1;
  out->edge = best_err.edge;
  out->line = best_err.line;
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

}  // namespace

// ---- plan object held by the ctx ------------------------------------------------------------
struct MsrcrPlan {
  llie_msrcr_params key;
  std::vector<ScalePlan> scales;
  std::vector<int> order;   // scale indices, level-C (six-tap) scales first
  // level-C maps the final kernel reads: one per stride-4 scale, then ONE for all stride-16 scales together
  // (up_bc_y_kernel: M = prod S_k^(weight_k / weight_ref), stored in the slot of the first stride-16 scale)
  struct FinalMap { int slot; float weight; bool four_tap; };
  std::vector<FinalMap> maps;
  std::vector<float> expo;  // per scale: exponent in M (stride-16 scales), 1 otherwise
  std::vector<int> accum;   // per scale: multiply into M (not the first stride-16 scale)
  double sum_e = 0;         // sum of the exponents inside M
  bool combined = false;    // false: every scale keeps its own map (M could leave the fp32 range)
  float* d_taps = nullptr;  // per-scale low-res taps
  std::vector<int> tap_off;
  UpWeights uw;
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

int msrcr_plan_detail(float sigma, float truncate, double* l1, double* edge, double* line, float* taps, int capacity,
                      int* count) {
  ScalePlan sp;
  LLIE_TRY(get_scale_plan(sigma, truncate, &sp));
  if (l1) *l1 = sp.l1;
  if (edge) *edge = sp.edge;
  if (line) *line = sp.line;
  if (count) *count = (int)sp.taps.size();
  if (taps) {
    LLIE_REQUIRE(capacity >= (int)sp.taps.size(), LLIE_E_ARG, "tap buffer too small: %d < %zu", capacity, sp.taps.size());
    std::copy(sp.taps.begin(), sp.taps.end(), taps);
  }
  return LLIE_OK;
}

static int ensure_plan(llie_ctx* ctx, const llie_msrcr_params& p, cudaStream_t st) {
  if (ctx->plan && memcmp(&ctx->plan->key, &p, sizeof(p)) == 0) return LLIE_OK;
  MsrcrPlan* pl = new MsrcrPlan;
  pl->key = p;
  std::vector<float> host;
  for (int k = 0; k < p.n_scales; ++k) {
    ScalePlan sp;
    int rc = get_scale_plan(p.sigma[k], p.truncate, &sp);
    if (rc != LLIE_OK) { delete pl; return rc; }
    pl->tap_off.push_back((int)host.size());
    host.insert(host.end(), sp.taps.begin(), sp.taps.end());
    host.resize((host.size() + 3) & ~size_t(3));
    pl->scales.push_back(std::move(sp));
  }
  for (int lv = 0; lv < 2; ++lv)
    for (int k = 0; k < p.n_scales; ++k)
      if (pl->scales[k].level == lv) pl->order.push_back(k);
  pl->expo.assign(p.n_scales, 1.f);
  pl->accum.assign(p.n_scales, 0);
  int kb0 = -1;
  float w_ref = 0.f;
  for (int k = 0; k < p.n_scales; ++k)
    if (pl->scales[k].level == 1) {
      if (kb0 < 0) kb0 = k;
      if (std::fabs(p.weight[k]) > std::fabs(w_ref)) w_ref = p.weight[k];
    }
  double sum_e = 0;
  for (int k = 0; k < p.n_scales; ++k)
    if (pl->scales[k].level == 1) sum_e += w_ref != 0.f ? std::fabs((double)p.weight[k] / w_ref) : 0.0;
  // M must stay a normal fp32 number: every surround lies in [in_bias, in_scale + in_bias]
  const bool combine = kb0 >= 0 && sum_e * std::fabs(std::log10((double)p.in_scale + p.in_bias)) <= 30.0 &&
                       sum_e * std::fabs(std::log10((double)p.in_bias)) <= 30.0;
  for (int k = 0; k < p.n_scales; ++k) {
    if (pl->scales[k].level == 0 || !combine) pl->maps.push_back({k, p.weight[k], false});
    else {
      pl->expo[k] = w_ref != 0.f ? p.weight[k] / w_ref : 0.f;
      pl->accum[k] = k != kb0;
    }
  }
  if (combine) {
    pl->maps.push_back({kb0, w_ref, true});
    pl->sum_e = sum_e;
  }
  pl->combined = combine;
  for (int ph = 0; ph < 4; ++ph) {
    double w[2 * kTU + 1];
    up_weights(ph, w);
    for (int t = 0; t <= 2 * kTU; ++t) pl->uw.w[ph][t] = (float)w[t];
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
  // the row-tiled path's captured graphs bake in pl->d_taps addresses: a new plan must change their key, or a later
  // call with the old parameters and buffers would replay launches that read the freed taps
  ++ctx->ws_generation;
  return LLIE_OK;
}

namespace {

inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// rows per strip so that the launch has at least `want` CTAs when the image allows it
int pick_strip(int rows, long long ctas_per_strip_row, int want, int lo, int hi, int mult) {
  int strips = std::max(1, ceil_div(want, std::max(1ll, ctas_per_strip_row)));
  int rps = ceil_div(rows, strips);
  rps = std::min(std::max(rps, lo), hi);
  return (rps + mult - 1) / mult * mult;
}

// Low-resolution separable Gaussian (planar): taps go by value when they fit the kernel-argument struct.
FirTaps make_fir_taps(const ScalePlan& sp, bool* fits) {
  FirTaps ft;
  memset(&ft, 0, sizeof(ft));
  *fits = fir8_steps(sp.radius_low) <= kFirParamTaps && (int)sp.taps.size() == 2 * sp.radius_low + 1;
  if (*fits)
    for (size_t i = 0; i < sp.taps.size(); ++i) ft.t[i] = sp.taps[i];
  return ft;
}
// The tap count of the default scales' filters (radius 16 at stride 4: 40 steps; 25 and 68 at stride 16: 64 and
// 144) selects a fully unrolled instantiation; anything else takes the dynamic loop.  `r` is the 4th argument.
template <class... A>
void launch_fir_x8(bool param, dim3 grid, size_t smem, cudaStream_t st, const float* in, float* out, const float* taps,
                   int r, int W, int rows_per_plane, int plane_rows, int row_lo, const FirTaps& ft) {
  const int steps = fir8_steps(r), total = (int)grid.x;  // grid.x arrives as the number of rows
  grid.x = (total + kFirXRows - 1) / kFirXRows;
  if (!param) {
    // long filters (taps in shared memory): two segment buffers of tile + steps floats can pass the 48 KB default;
    // beyond the opt-in limit the launch itself fails and LLIE_LAUNCH reports it
    if (smem > 48 * 1024) cudaFuncSetAttribute(fir_x8_kernel<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    fir_x8_kernel<false, 0><<<grid, 128, smem, st>>>(in, out, taps, r, W, rows_per_plane, plane_rows, row_lo, total, ft);
  }
  else if (steps == 40) fir_x8_kernel<true, 40><<<grid, 128, smem, st>>>(in, out, taps, r, W, rows_per_plane, plane_rows, row_lo, total, ft);
  else if (steps == 64) fir_x8_kernel<true, 64><<<grid, 128, smem, st>>>(in, out, taps, r, W, rows_per_plane, plane_rows, row_lo, total, ft);
  else if (steps == 144) fir_x8_kernel<true, 144><<<grid, 128, smem, st>>>(in, out, taps, r, W, rows_per_plane, plane_rows, row_lo, total, ft);
  else fir_x8_kernel<true, 0><<<grid, 128, smem, st>>>(in, out, taps, r, W, rows_per_plane, plane_rows, row_lo, total, ft);
}
template <class... A>
void launch_fir_y8(bool param, dim3 grid, size_t smem, cudaStream_t st, const float* in, float* out, const float* taps,
                   int r, A... a) {
  const int steps = fir8_steps(r);
  if (!param) fir_y8_kernel<false, 0><<<grid, 128, smem, st>>>(in, out, taps, r, a...);
  else if (steps == 40) fir_y8_kernel<true, 40><<<grid, 128, smem, st>>>(in, out, taps, r, a...);
  else if (steps == 64) fir_y8_kernel<true, 64><<<grid, 128, smem, st>>>(in, out, taps, r, a...);
  else if (steps == 144) fir_y8_kernel<true, 144><<<grid, 128, smem, st>>>(in, out, taps, r, a...);
  else fir_y8_kernel<true, 0><<<grid, 128, smem, st>>>(in, out, taps, r, a...);
}

template <int N6, int N4>
void launch_final(dim3 grid, cudaStream_t st, const void* x, void* out, int iot, ImgView v, int io, int W, int in_row0,
                  int row_begin, int row_end, int rps, const FinalParams& fp, bool eq) {
  // eq: equal weights over the maps of this launch and values bounded so that their product stays a normal fp32
  // number (decided by final_passes, which knows the exponents inside the combined stride-16 map)
  for (int i = 1; i < N6 + N4; ++i) eq = eq && fp.w2[i] == fp.w2[0];
  if constexpr (N6 == 1 && N4 == 1) {
    // the default call (equal weights, one pass, vector I/O): layout-specialised instantiations
    if (eq && !fp.acc_in && !fp.acc_out && (io == kIoNhwcVec || io == kIoNchwVec)) {
      const bool nhwc = io == kIoNhwcVec;
      if (iot == kIoU16) {
        if (nhwc) msrcr_final_kernel<1, 1, uint16_t, true, kIoNhwcVec><<<grid, 128, 0, st>>>((const uint16_t*)x, (uint16_t*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
        else msrcr_final_kernel<1, 1, uint16_t, true, kIoNchwVec><<<grid, 128, 0, st>>>((const uint16_t*)x, (uint16_t*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
      } else if (iot == kIoU8) {
        if (nhwc) msrcr_final_kernel<1, 1, uint8_t, true, kIoNhwcVec><<<grid, 128, 0, st>>>((const uint8_t*)x, (uint8_t*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
        else msrcr_final_kernel<1, 1, uint8_t, true, kIoNchwVec><<<grid, 128, 0, st>>>((const uint8_t*)x, (uint8_t*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
      } else {
        if (nhwc) msrcr_final_kernel<1, 1, float, true, kIoNhwcVec><<<grid, 128, 0, st>>>((const float*)x, (float*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
        else msrcr_final_kernel<1, 1, float, true, kIoNchwVec><<<grid, 128, 0, st>>>((const float*)x, (float*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
      }
      return;
    }
  }
  if (iot == kIoU16) {
    if (eq) msrcr_final_kernel<N6, N4, uint16_t, true><<<grid, 128, 0, st>>>((const uint16_t*)x, (uint16_t*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
    else msrcr_final_kernel<N6, N4, uint16_t, false><<<grid, 128, 0, st>>>((const uint16_t*)x, (uint16_t*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
  } else if (iot == kIoU8) {
    if (eq) msrcr_final_kernel<N6, N4, uint8_t, true><<<grid, 128, 0, st>>>((const uint8_t*)x, (uint8_t*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
    else msrcr_final_kernel<N6, N4, uint8_t, false><<<grid, 128, 0, st>>>((const uint8_t*)x, (uint8_t*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
  } else {
    if (eq) msrcr_final_kernel<N6, N4, float, true><<<grid, 128, 0, st>>>((const float*)x, (float*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
    else msrcr_final_kernel<N6, N4, float, false><<<grid, 128, 0, st>>>((const float*)x, (float*)out, v, io, W, in_row0, row_begin, row_end, rps, fp);
  }
}


// K5 launches: rows [row_begin, row_end) of n images whose pointers start at image row in_row0.
int final_passes(llie_ctx* ctx, cudaStream_t st, const MsrcrPlan* pl, const llie_msrcr_params& p, const void* in,
                 void* out, const ImgView& v, bool vec, bool nhwc, int n, int h_view, int w, int in_row0, int row_begin,
                 int row_end, int hC, int wC, const float* dSC, size_t c_elems, float* dAcc, int iot = kIoF32) {
  (void)h_view;
  const int io = vec ? (nhwc ? kIoNhwcVec : kIoNchwVec) : kIoScalar;
  const int cx = ceil_div(w, 4 * kFinalCells);
  const int rps = pick_strip(row_end - row_begin, (long long)cx * n, 16 * ctx->sm_count, 16, 64, 4);
  dim3 grid(cx, ceil_div(row_end - row_begin, rps), n);
  FinalParams fp;
  memset(&fp, 0, sizeof(fp));
  fp.hC = hC; fp.wC = wC;
  const double ln2 = 0.6931471805599453;
  double wsum = 0;
  for (int k = 0; k < p.n_scales; ++k) wsum += p.weight[k];
  fp.wsum2 = (float)(ln2 * wsum);
  fp.beta_ln_alpha = (float)(p.beta * std::log((double)p.alpha));
  fp.beta_ln2 = (float)(p.beta * ln2);
  fp.gain255 = p.gain / 255.f;
  fp.offset255 = p.offset / 255.f;
  fp.in_scale = p.in_scale / io_levels(iot); fp.in_bias = p.in_bias;  // integer images are read as raw 0..255 / 0..65535
  fp.no_clip = (iot == kIoF32 && p.finish != LLIE_FINISH_NONE) ? 1 : 0;
  const int nm = (int)pl->maps.size();
  // one MUFU.LG2 of the PRODUCT of a launch's maps is only taken while that product provably stays a normal fp32
  // number: every surround lies in [in_bias, in_scale + in_bias], the combined map carries sum_e of them
  const double lmax = std::log10((double)p.in_scale + p.in_bias), lmin = std::log10((double)p.in_bias);
  auto product_ok = [&](int k0, int cnt) {
    double e = 0;
    for (int i = 0; i < cnt; ++i) e += pl->maps[k0 + i].four_tap ? pl->sum_e : 1.0;
    return e * std::max(lmax, 0.0) <= 36.0 && e * std::min(lmin, 0.0) >= -36.0;
  };
  if (nm == 2 && !pl->maps[0].four_tap && pl->maps[1].four_tap) {
    // the default shape (one stride-4 scale + the combined stride-16 map): 6-tap + 4-tap upsample in one pass
    for (int i = 0; i < 2; ++i) {
      fp.map[i] = dSC + (size_t)pl->maps[i].slot * c_elems;
      fp.w2[i] = (float)(ln2 * pl->maps[i].weight);
    }
    LLIE_LAUNCH(ctx, st, "msrcr_final_kernel", (launch_final<1, 1>(grid, st, in, out, iot, v, io, w, in_row0, row_begin, row_end, rps, fp, product_ok(0, 2))));
  } else {
    for (int k0 = 0; k0 < nm; k0 += 3) {
      const int cnt = std::min(3, nm - k0);
      for (int i = 0; i < cnt; ++i) {
        fp.map[i] = dSC + (size_t)pl->maps[k0 + i].slot * c_elems;
        fp.w2[i] = (float)(ln2 * pl->maps[k0 + i].weight);
      }
      fp.acc_in = k0 > 0 ? (const float*)dAcc : nullptr;
      fp.acc_out = k0 + cnt < nm ? dAcc : nullptr;
      if (cnt == 1) LLIE_LAUNCH(ctx, st, "msrcr_final_kernel", (launch_final<1, 0>(grid, st, in, out, iot, v, io, w, in_row0, row_begin, row_end, rps, fp, product_ok(k0, cnt))));
      if (cnt == 2) LLIE_LAUNCH(ctx, st, "msrcr_final_kernel", (launch_final<2, 0>(grid, st, in, out, iot, v, io, w, in_row0, row_begin, row_end, rps, fp, product_ok(k0, cnt))));
      if (cnt == 3) LLIE_LAUNCH(ctx, st, "msrcr_final_kernel", (launch_final<3, 0>(grid, st, in, out, iot, v, io, w, in_row0, row_begin, row_end, rps, fp, product_ok(k0, cnt))));
    }
  }
  return LLIE_OK;
}

// B -> C upsample (y pass) of the stride-16 scales listed in `ks`: one launch per pair writes / multiplies their
// (powered) product into the combined map, or -- without combination -- every scale into its own slot.
int launch_up_bc_y(llie_ctx* ctx, cudaStream_t st, const MsrcrPlan* pl, const std::vector<int>& ks, const float* t2_base,
                   size_t t2_elems, float* sc_base, size_t c_elems, int hB, int hC, int wC, int c0, int c1, int planes) {
  dim3 uy(ceil_div(wC, 128), up_bc_y_chunks(c0, c1), planes);
  if (!pl->combined) {
    for (size_t i = 0; i < ks.size(); ++i) {
      UpSrc u{{t2_base + i * t2_elems, nullptr}, {1.f, 1.f}};
      LLIE_LAUNCH(ctx, st, "up_bc_y_kernel", (up_bc_y_kernel<1><<<uy, 128, 0, st>>>(u, sc_base + (size_t)ks[i] * c_elems, hB, hC, wC, c0, c1, pl->uw, 0)));
    }
    return LLIE_OK;
  }
  float* dst = sc_base + (size_t)pl->maps.back().slot * c_elems;
  for (size_t i = 0; i < ks.size(); i += 2) {
    if (i + 1 < ks.size()) {
      UpSrc u{{t2_base + i * t2_elems, t2_base + (i + 1) * t2_elems}, {pl->expo[ks[i]], pl->expo[ks[i + 1]]}};
      LLIE_LAUNCH(ctx, st, "up_bc_y_kernel", (up_bc_y_kernel<2><<<uy, 128, 0, st>>>(u, dst, hB, hC, wC, c0, c1, pl->uw, i > 0)));
    } else {
      UpSrc u{{t2_base + i * t2_elems, nullptr}, {pl->expo[ks[i]], 1.f}};
      LLIE_LAUNCH(ctx, st, "up_bc_y_kernel", (up_bc_y_kernel<1><<<uy, 128, 0, st>>>(u, dst, hB, hC, wC, c0, c1, pl->uw, i > 0)));
    }
  }
  return LLIE_OK;
}

}  // namespace

int msrcr_pyramid_run(llie_ctx* ctx, const float* in, float* out, int n, int h, int w, int layout,
                      const llie_msrcr_params& p, cudaStream_t st) {
  return msrcr_pyramid_run_io(ctx, in, out, false, n, h, w, layout, p, st);
}

// u8: in / out are uint8_t images (value = byte / 255 on the way in, rint(255 * value) on the way out), converted
// inside the first and last kernels -- no extra pass over the image.
int msrcr_pyramid_run_io(llie_ctx* ctx, const void* in, void* out, int iot, int n, int h, int w, int layout,
                         const llie_msrcr_params& p, cudaStream_t st) {
  LLIE_TRY(ensure_plan(ctx, p, st));
  MsrcrPlan* pl = ctx->plan;
  const int hC = level_c_size(h), wC = level_c_size(w), hB = level_b_size(h), wB = level_b_size(w);
  const int planes = n * 3;
  const bool nhwc = layout == LLIE_NHWC;
  const int RW = nhwc ? 3 * w : w, CS = nhwc ? 3 : 1, planes_v = nhwc ? n : planes;
  int n_b = 0;
  for (auto& s : pl->scales) n_b += s.level;
  const size_t c_elems = (size_t)planes * hC * wC, b_elems = (size_t)planes * hB * wB;
  void *dV, *dC, *dCt, *dSC, *dTB = nullptr, *dB = nullptr, *dBt = nullptr, *dSB = nullptr, *dT2 = nullptr, *dAcc = nullptr;
  LLIE_TRY(ctx->workspace("msrcr.V", (size_t)planes_v * hC * RW * 4, &dV));
  LLIE_TRY(ctx->workspace("msrcr.C", c_elems * 4, &dC));
  LLIE_TRY(ctx->workspace("msrcr.C.tmp", c_elems * 4, &dCt));
  LLIE_TRY(ctx->workspace("msrcr.SC", c_elems * 4 * p.n_scales, &dSC));
  if (n_b) {
    LLIE_TRY(ctx->workspace("msrcr.TB", (size_t)planes * hC * wB * 4, &dTB));
    LLIE_TRY(ctx->workspace("msrcr.B", b_elems * 4, &dB));
    LLIE_TRY(ctx->workspace("msrcr.B.tmp", b_elems * 4, &dBt));
    LLIE_TRY(ctx->workspace("msrcr.SB", b_elems * 4, &dSB));
    LLIE_TRY(ctx->workspace("msrcr.T2", (size_t)n_b * planes * hB * wC * 4, &dT2));
  }
  if (pl->maps.size() > 3) LLIE_TRY(ctx->workspace("msrcr.acc", (size_t)n * h * w * 3 * 4, &dAcc));

  const ImgView v = make_view(layout, 3, h, w);
  const bool vec = (w % 4 == 0) && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) % (iot == kIoU8 ? 4 : (iot == kIoU16 ? 8 : 16)) == 0);
  const int want = 8 * ctx->sm_count;
  const float in_scale = p.in_scale / io_levels(iot);

  {  // K1: rows
    const long long plane_stride = nhwc ? (long long)h * RW : (long long)h * w;
    if (vec) {
      const int cx = ceil_div(RW, 512);
      const int rps = pick_strip(hC, (long long)cx * planes_v, want, 16, 64, 1);
      dim3 grid(cx, ceil_div(hC, rps), planes_v);
      if (iot == kIoU16)
        LLIE_LAUNCH(ctx, st, "down4_rows_kernel",
                    (down4_rows_kernel<4, uint16_t><<<grid, 128, 0, st>>>((const uint16_t*)in, plane_stride, RW, h, 0, h, (float*)dV, hC, 0, hC, rps)));
      else if (iot == kIoU8)
        LLIE_LAUNCH(ctx, st, "down4_rows_kernel",
                    (down4_rows_kernel<4, uint8_t><<<grid, 128, 0, st>>>((const uint8_t*)in, plane_stride, RW, h, 0, h, (float*)dV, hC, 0, hC, rps)));
      else
        LLIE_LAUNCH(ctx, st, "down4_rows_kernel",
                    (down4_rows_kernel<4, float><<<grid, 128, 0, st>>>((const float*)in, plane_stride, RW, h, 0, h, (float*)dV, hC, 0, hC, rps)));
    } else {
      const int cx = ceil_div(RW, 128);
      const int rps = pick_strip(hC, (long long)cx * planes_v, want, 16, 64, 1);
      dim3 grid(cx, ceil_div(hC, rps), planes_v);
      if (iot == kIoU16)
        LLIE_LAUNCH(ctx, st, "down4_rows_kernel",
                    (down4_rows_kernel<1, uint16_t><<<grid, 128, 0, st>>>((const uint16_t*)in, plane_stride, RW, h, 0, h, (float*)dV, hC, 0, hC, rps)));
      else if (iot == kIoU8)
        LLIE_LAUNCH(ctx, st, "down4_rows_kernel",
                    (down4_rows_kernel<1, uint8_t><<<grid, 128, 0, st>>>((const uint8_t*)in, plane_stride, RW, h, 0, h, (float*)dV, hC, 0, hC, rps)));
      else
        LLIE_LAUNCH(ctx, st, "down4_rows_kernel",
                    (down4_rows_kernel<1, float><<<grid, 128, 0, st>>>((const float*)in, plane_stride, RW, h, 0, h, (float*)dV, hC, 0, hC, rps)));
    }
  }
  {  // K1b: columns
    dim3 grid(ceil_div(wC, 128), ceil_div(hC, kColsRows), planes_v);
    if (nhwc)
      LLIE_LAUNCH(ctx, st, "down4_cols_kernel",
                  down4_cols4_kernel<<<dim3(ceil_div(wC, kCols4Out), grid.y, grid.z), 128, 0, st>>>((const float*)dV, (float*)dC, RW, w, hC, wC, 0, hC, in_scale, p.in_bias));
    else
      LLIE_LAUNCH(ctx, st, "down4_cols_kernel",
                  down4_cols_kernel<1><<<grid, 128, 0, st>>>((const float*)dV, (float*)dC, RW, w, hC, wC, 0, hC, in_scale, p.in_bias));
  }
  // The stride-16 branch (decimation, the stride-16 FIRs, both upsamplers: ten small, latency-bound launches on
  // 1/16 .. 1/256 of the pixels) and the stride-4 FIRs (issue-bound) both depend on level C only: they run on two
  // streams, forked and joined by events, so the small launches fill in underneath the large ones.
  int n_c = 0;
  for (auto& sc0 : pl->scales) n_c += sc0.level == 0;
  cudaStream_t sb = st;
  if (n_b && n_c) {
    LLIE_TRY(ensure_aux_stream(ctx));
    sb = ctx->aux_stream;
    LLIE_CUDA_OK(cudaEventRecord(ctx->aux_ev[0], st));
    LLIE_CUDA_OK(cudaStreamWaitEvent(sb, ctx->aux_ev[0], 0));
  }
  if (n_b) {
    LLIE_LAUNCH(ctx, sb, "decim4_x_kernel", decim4_x_kernel<<<dim3(ceil_div(wB, kDecXTile), hC, planes), 128, 0, sb>>>((const float*)dC, (float*)dTB, hC, wC, wB, 0, hC));
    LLIE_LAUNCH(ctx, sb, "decim4_y_kernel", decim4_y_kernel<<<dim3(ceil_div(wB, 128), ceil_div(hB, kDecQ), planes), 128, 0, sb>>>((const float*)dTB, (float*)dB, hC, hB, wB, 0, hB));
  }
  (void)CS;
  std::vector<int> b_scales;  // stride-16 scales in T2 order
  const size_t t2_elems = (size_t)planes * hB * wC;
  for (int k = 0; k < p.n_scales; ++k) {
    const ScalePlan& sp = pl->scales[k];
    const float* taps = pl->d_taps + pl->tap_off[k];
    const int r = sp.radius_low;
    const size_t smx = fir_x8_smem(r), smy = (size_t)fir8_steps(r) * sizeof(float);
    bool tp;
    const FirTaps ft = make_fir_taps(sp, &tp);
    float* sc = (float*)dSC + (size_t)k * c_elems;
    if (sp.level == 0) {
      dim3 gx((unsigned)((long long)planes * hC), ceil_div(wC, kFirX8Tile)), gy(ceil_div(wC, 128), ceil_div(hC, kFir8), planes);
      LLIE_LAUNCH(ctx, st, "fir_x_kernel", (launch_fir_x8(tp, gx, smx, st, (const float*)dC, (float*)dCt, taps, r, wC, planes * hC, 0, 0, ft)));
      LLIE_LAUNCH(ctx, st, "fir_y_kernel", (launch_fir_y8(tp, gy, smy, st, (const float*)dCt, sc, taps, r, hC, wC, 0, hC, 0, hC, 0, hC, ft)));
    } else {
      dim3 gx((unsigned)((long long)planes * hB), ceil_div(wB, kFirX8Tile)), gy(ceil_div(wB, 128), ceil_div(hB, kFir8), planes);
      LLIE_LAUNCH(ctx, sb, "fir_x_kernel", (launch_fir_x8(tp, gx, smx, sb, (const float*)dB, (float*)dBt, taps, r, wB, planes * hB, 0, 0, ft)));
      LLIE_LAUNCH(ctx, sb, "fir_y_kernel", (launch_fir_y8(tp, gy, smy, sb, (const float*)dBt, (float*)dSB, taps, r, hB, wB, 0, hB, 0, hB, 0, hB, ft)));
      dim3 ux(ceil_div((long long)hB * up_bc_x_groups(wC), 256), 1, planes);
      LLIE_LAUNCH(ctx, sb, "up_bc_x_kernel", up_bc_x_kernel<<<ux, 256, 0, sb>>>((const float*)dSB, (float*)dT2 + b_scales.size() * t2_elems, hB, wB, wC, 0, hB, pl->uw));
      b_scales.push_back(k);
    }
  }
  if (!b_scales.empty())
    LLIE_TRY(launch_up_bc_y(ctx, sb, pl, b_scales, (const float*)dT2, t2_elems, (float*)dSC, c_elems, hB, hC, wC, 0, hC, planes));
  if (sb != st) {
    LLIE_CUDA_OK(cudaEventRecord(ctx->aux_ev[1], sb));
    LLIE_CUDA_OK(cudaStreamWaitEvent(st, ctx->aux_ev[1], 0));
  }
  LLIE_TRY(final_passes(ctx, st, pl, p, in, out, v, vec, nhwc, n, h, w, 0, 0, h, hC, wC, (const float*)dSC, c_elems,
                        (float*)dAcc, iot));
  return LLIE_OK;
}


// ---- row-tiled variant (config 5) -----------------------------------------------------------------
// One image of H x W pixels is split into horizontal shards, one per device.  Every level of the
// cascade keeps its GLOBAL row indexing on every device (buffers are sized for the whole image; only
// the rows a device needs are ever touched), so the single-device kernels run unchanged on row
// ranges.  What crosses devices:
//   1. 16 raw image rows from each vertical neighbour (the stride-4 prefilter's reach), copied into
//      small "band" buffers so that the boundary V rows are computed locally;
//   2. stride-4 level-C rows: every device computes the C rows centred on its own pixels and then
//      fetches the C rows its large-sigma surrounds reach (about 4*sigma_max/4 rows each way) from
//      whichever devices own them.  For 16384 x 16384 and sigma = 250 that is ~15 MB per side
//      instead of the 271 MB of full-resolution row-blurred halos.
// Everything after the C-level gather (B level, low-resolution FIRs, upsampling, epilogue) is local
// and recomputed redundantly on the overlap, which is 1/16 .. 1/256 of the pixel count.
// Copies are cudaMemcpyPeerAsync over NVLink (plain device-to-device copies when the shards are
// "virtual devices" on one GPU); there is no NCCL and no collective.
namespace {

struct TileGeom {
  int R0, R1;            // image rows of the shard
  int own_lo, own_hi;    // C rows this device computes from its own pixels
  int s_lo, s_hi;        // C-resolution surround rows the final kernel reads
  int c0_lo, c0_hi;      // C rows needed by the stride-4 (level C) scales
  int t_lo, t_hi;        // B-resolution surround rows (SB / T2) needed by the upsampler
  int bb_lo, bb_hi;      // B rows needed by the stride-16 FIRs
  int c1_lo, c1_hi;      // C rows needed to build those B rows
  int cn_lo, cn_hi;      // union: all C rows this device must hold
};

inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

int up_bc_j0(int jc) {  // mirrors up_bc_phase()
  const int d = jc - kPC + 4, ph = d & 3;
  return (d >> 2) - 1 + kPB + (ph >= 2 ? 1 : 0);
}

TileGeom tile_geometry(int d, int n_dev, const int* row0, const int* rows, int H, int r0_max, int rb_max, bool any_b) {
  const int hC = level_c_size(H), hB = level_b_size(H);
  TileGeom g;
  g.R0 = row0[d];
  g.R1 = row0[d] + rows[d];
  g.own_lo = d == 0 ? 0 : g.R0 / 4 + kPC;
  g.own_hi = d == n_dev - 1 ? hC : g.R1 / 4 + kPC;
  g.s_lo = clampi(g.R0 / 4 + kPC - 2, 0, hC);
  g.s_hi = clampi((g.R1 - 1) / 4 + kPC + 5, 0, hC);
  g.c0_lo = clampi(g.s_lo - r0_max, 0, hC);
  g.c0_hi = clampi(g.s_hi + r0_max, 0, hC);
  g.cn_lo = g.c0_lo;
  g.cn_hi = g.c0_hi;
  g.t_lo = g.t_hi = g.bb_lo = g.bb_hi = g.c1_lo = g.c1_hi = 0;
  if (any_b) {
    g.t_lo = clampi(up_bc_j0(g.s_lo) - kTU, 0, hB);
    g.t_hi = clampi(up_bc_j0(g.s_hi - 1) + kTU + 1, 0, hB);
    g.bb_lo = clampi(g.t_lo - rb_max, 0, hB);
    g.bb_hi = clampi(g.t_hi + rb_max, 0, hB);
    g.c1_lo = clampi(4 * (g.bb_lo - kPB) + kPC - kRD, 0, hC - 1);
    g.c1_hi = clampi(4 * (g.bb_hi - 1 - kPB) + kPC + kRD, 0, hC - 1) + 1;
    g.cn_lo = std::min(g.cn_lo, g.c1_lo);
    g.cn_hi = std::max(g.cn_hi, g.c1_hi);
  }
  g.cn_lo = std::min(g.cn_lo, g.own_lo);
  g.cn_hi = std::max(g.cn_hi, g.own_hi);
  return g;
}

}  // namespace


namespace {

struct TileBufs {
  float *V, *C, *Ct, *SC, *TB, *B, *Bt, *SB, *T2, *band_top, *band_bot;
  void *fin_part = nullptr, *fin_gather = nullptr, *fin_li = nullptr, *fin_state = nullptr;  // global-statistics finisher
};

constexpr int kBandTopRows = 48, kBandBotRows = 64, kHaloRaw = 16;

int copy_dev(void* dst, int dst_dev, const void* src, int src_dev, size_t bytes, cudaStream_t st) {
  if (dst_dev == src_dev) LLIE_CUDA_OK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, st));
  else LLIE_CUDA_OK(cudaMemcpyPeerAsync(dst, dst_dev, src, src_dev, bytes, st));
  return LLIE_OK;
}

// Phase A on device d: raw-row bands from the neighbours, V, and the C rows centred on own pixels.
int tiled_phase_a_body(llie_ctx* ctx, cudaStream_t st, int d, int n_dev, llie_ctx* const* ctxs, const float* const* in,
                       const int* rows, const TileGeom& g, const TileBufs& b, int H, int W, const llie_msrcr_params& p) {
  const int RW = 3 * W, hC = level_c_size(H), wC = level_c_size(W);
  const bool vec = (W % 4 == 0) && (reinterpret_cast<uintptr_t>(in[d]) % 16 == 0);
  const bool first = d == 0, last = d == n_dev - 1;
  const size_t row_bytes = (size_t)RW * sizeof(float);
  if (!first) {  // rows [R0-16, R0) from the device above, [R0, R0+32) own
    LLIE_TRY(copy_dev(b.band_top, ctx->device, in[d - 1] + (size_t)(rows[d - 1] - kHaloRaw) * RW, ctxs[d - 1]->device,
                      kHaloRaw * row_bytes, st));
    LLIE_TRY(copy_dev(b.band_top + (size_t)kHaloRaw * RW, ctx->device, in[d], ctx->device,
                      (kBandTopRows - kHaloRaw) * row_bytes, st));
  }
  if (!last) {  // rows [R1-48, R1) own, [R1, R1+16) from the device below
    LLIE_TRY(copy_dev(b.band_bot, ctx->device, in[d] + (size_t)(rows[d] - (kBandBotRows - kHaloRaw)) * RW, ctx->device,
                      (kBandBotRows - kHaloRaw) * row_bytes, st));
    LLIE_TRY(copy_dev(b.band_bot + (size_t)(kBandBotRows - kHaloRaw) * RW, ctx->device, in[d + 1], ctxs[d + 1]->device,
                      kHaloRaw * row_bytes, st));
  }
  const int vi_lo = first ? 0 : g.R0 / 4 + kPC + 4;                  // V rows whose 33-tap window stays inside the shard
  const int vi_hi = last ? hC : (g.R1 - 17) / 4 + kPC + 1;           // (or clamps at a true image border)
  auto rows_launch = [&](const float* x, int in_row0, int in_rows, int v0, int v1, int want) -> int {
    if (v0 >= v1) return LLIE_OK;
    if (vec) {
      const int cx = ceil_div(RW, 512);
      const int rps = pick_strip(v1 - v0, cx, want, 4, 64, 1);
      dim3 grid(cx, ceil_div(v1 - v0, rps), 1);
      LLIE_LAUNCH(ctx, st, "down4_rows_kernel",
                  (down4_rows_kernel<4, float><<<grid, 128, 0, st>>>(x, 0, RW, H, in_row0, in_rows, b.V, hC, v0, v1, rps)));
    } else {
      const int cx = ceil_div(RW, 128);
      const int rps = pick_strip(v1 - v0, cx, want, 4, 64, 1);
      dim3 grid(cx, ceil_div(v1 - v0, rps), 1);
      LLIE_LAUNCH(ctx, st, "down4_rows_kernel",
                  (down4_rows_kernel<1, float><<<grid, 128, 0, st>>>(x, 0, RW, H, in_row0, in_rows, b.V, hC, v0, v1, rps)));
    }
    return LLIE_OK;
  };
  if (!first) LLIE_TRY(rows_launch(b.band_top, g.R0 - kHaloRaw, kBandTopRows, g.own_lo, vi_lo, 1));
  if (!last) LLIE_TRY(rows_launch(b.band_bot, g.R1 - (kBandBotRows - kHaloRaw), kBandBotRows, vi_hi, g.own_hi, 1));
  LLIE_TRY(rows_launch(in[d], g.R0, rows[d], vi_lo, vi_hi, 8 * ctx->sm_count));
  dim3 grid(ceil_div(wC, kCols4Out), ceil_div(g.own_hi - g.own_lo, kColsRows), 1);
  LLIE_LAUNCH(ctx, st, "down4_cols_kernel",
              down4_cols4_kernel<<<grid, 128, 0, st>>>(b.V, b.C, RW, W, hC, wC, g.own_lo, g.own_hi, p.in_scale, p.in_bias));
  return LLIE_OK;
}

uint64_t tile_key(llie_ctx* ctx, int d, int n_dev, const void* in_d, const void* in_up, const void* in_dn, const void* out_d,
                  int rows_up, const TileGeom& g, int H, int W, const llie_msrcr_params& p) {
  struct { uint64_t gen; const void *a, *b, *c, *o; int d, n, H, W, ru; TileGeom g; llie_msrcr_params p; } k;
  memset(&k, 0, sizeof(k));
  k.gen = ctx->ws_generation; k.a = in_d; k.b = in_up; k.c = in_dn; k.o = out_d; k.d = d; k.n = n_dev; k.H = H; k.W = W;
  k.ru = rows_up; k.g = g; k.p = p;
  return fnv1a(&k, sizeof(k)) | 1ull;
}

int tiled_phase_a(llie_ctx* ctx, int d, int n_dev, llie_ctx* const* ctxs, const float* const* in, const int* rows,
                  const TileGeom& g, const TileBufs& b, int H, int W, const llie_msrcr_params& p) {
  const uint64_t key = tile_key(ctx, d, n_dev, in[d], d > 0 ? in[d - 1] : nullptr, d + 1 < n_dev ? in[d + 1] : nullptr, nullptr,
                                d > 0 ? rows[d - 1] : 0, g, H, W, p);
  return run_as_graph(ctx, ctx->g_tile[0], key, ctx->stream, [&](cudaStream_t st) {
    return tiled_phase_a_body(ctx, st, d, n_dev, ctxs, in, rows, g, b, H, W, p);
  });
}

// Phase B on device d: fetch the foreign C rows, then everything downstream, locally.
int tiled_phase_b(llie_ctx* ctx, int d, int n_dev, llie_ctx* const* ctxs, const float* const* in, float* const* out,
                  const std::vector<TileGeom>& geo, const std::vector<TileBufs>& bufs, const std::vector<cudaEvent_t>& ev_c,
                  int H, int W, const llie_msrcr_params& p) {
  cudaStream_t st = ctx->stream;
  const TileGeom& g = geo[d];
  const TileBufs& b = bufs[d];
  const MsrcrPlan* pl = ctx->plan;
  const int hC = level_c_size(H), wC = level_c_size(W), hB = level_b_size(H), wB = level_b_size(W);
  const size_t c_elems = (size_t)3 * hC * wC;
  // Foreign C rows.  The rows the stride-4 scales reach (c0 range: a few rows beyond the shard) are fetched on the
  // compute stream; the far rows only the stride-16 branch needs travel on the ctx's copy stream underneath the
  // stride-4 FIRs (order the work so that the copies overlap it) and are joined before decim4_x.
  cudaStream_t far = ctx->sh_h2d ? ctx->sh_h2d : st;
  bool far_used = false;
  auto fetch = [&](int e, int lo, int hi, cudaStream_t s2) -> int {
    if (lo >= hi) return LLIE_OK;
    for (int pln = 0; pln < 3; ++pln) {
      const size_t off = ((size_t)pln * hC + lo) * wC;
      LLIE_TRY(copy_dev(b.C + off, ctx->device, bufs[e].C + off, ctxs[e]->device, (size_t)(hi - lo) * wC * sizeof(float), s2));
    }
    return LLIE_OK;
  };
  for (int e = 0; e < n_dev; ++e) {
    if (e == d) continue;
    const int lo = std::max(g.cn_lo, geo[e].own_lo), hi = std::min(g.cn_hi, geo[e].own_hi);
    if (lo >= hi) continue;
    const int nlo = std::max(lo, g.c0_lo), nhi = std::min(hi, g.c0_hi);  // near part (may be empty)
    LLIE_CUDA_OK(cudaStreamWaitEvent(st, ev_c[e], 0));
    if (nlo < nhi) {
      LLIE_TRY(fetch(e, nlo, nhi, st));
      if (far != st && (lo < nlo || nhi < hi)) {
        LLIE_CUDA_OK(cudaStreamWaitEvent(far, ev_c[e], 0));
        far_used = true;
      }
      LLIE_TRY(fetch(e, lo, nlo, far));
      LLIE_TRY(fetch(e, nhi, hi, far));
    } else {
      if (far != st) {
        LLIE_CUDA_OK(cudaStreamWaitEvent(far, ev_c[e], 0));
        far_used = true;
      }
      LLIE_TRY(fetch(e, lo, hi, far));
    }
  }
  if (far_used) LLIE_CUDA_OK(cudaEventRecord(ctx->sh_ev[0][0], far));
  // Stride-4 scales first (plain launches): they need the near rows only, and the far copies run underneath.  The
  // rest -- decimation to stride 16, the stride-16 scales, both upsamplers and the final kernel: ~11 dependent
  // launches -- is one CUDA graph.
  const size_t t2_elems = (size_t)3 * hB * wC;
  auto scale_launches = [&](cudaStream_t s2, int level, std::vector<int>* b_scales) -> int {
    for (int k = 0; k < p.n_scales; ++k) {
      const ScalePlan& sp = pl->scales[k];
      if (sp.level != level) continue;
      const float* taps = pl->d_taps + pl->tap_off[k];
      const int r = sp.radius_low;
      const size_t smx = fir_x8_smem(r), smy = (size_t)fir8_steps(r) * sizeof(float);
      bool tp;
      const FirTaps ft = make_fir_taps(sp, &tp);
      float* sc = b.SC + (size_t)k * c_elems;
      if (sp.level == 0) {
        const int x0 = clampi(g.s_lo - r, 0, hC), x1 = clampi(g.s_hi + r, 0, hC);
        dim3 gx(3 * (x1 - x0), ceil_div(wC, kFirX8Tile)), gy(ceil_div(wC, 128), ceil_div(g.s_hi - g.s_lo, kFir8), 3);
        LLIE_LAUNCH(ctx, s2, "fir_x_kernel", (launch_fir_x8(tp, gx, smx, s2, b.C, b.Ct, taps, r, wC, x1 - x0, hC, x0, ft)));
        LLIE_LAUNCH(ctx, s2, "fir_y_kernel", (launch_fir_y8(tp, gy, smy, s2, b.Ct, sc, taps, r, hC, wC, 0, hC, g.s_lo, g.s_hi - g.s_lo, g.s_lo, hC, ft)));
      } else {
        const int x0 = clampi(g.t_lo - r, 0, hB), x1 = clampi(g.t_hi + r, 0, hB);
        dim3 gx(3 * (x1 - x0), ceil_div(wB, kFirX8Tile)), gy(ceil_div(wB, 128), ceil_div(g.t_hi - g.t_lo, kFir8), 3);
        LLIE_LAUNCH(ctx, s2, "fir_x_kernel", (launch_fir_x8(tp, gx, smx, s2, b.B, b.Bt, taps, r, wB, x1 - x0, hB, x0, ft)));
        LLIE_LAUNCH(ctx, s2, "fir_y_kernel", (launch_fir_y8(tp, gy, smy, s2, b.Bt, b.SB, taps, r, hB, wB, 0, hB, g.t_lo, g.t_hi - g.t_lo, g.t_lo, hB, ft)));
        dim3 ux(ceil_div((long long)(g.t_hi - g.t_lo) * up_bc_x_groups(wC), 256), 1, 3);
        LLIE_LAUNCH(ctx, s2, "up_bc_x_kernel", up_bc_x_kernel<<<ux, 256, 0, s2>>>(b.SB, b.T2 + b_scales->size() * t2_elems, hB, wB, wC, g.t_lo, g.t_hi, pl->uw));
        b_scales->push_back(k);
      }
    }
    return LLIE_OK;
  };
  LLIE_TRY(scale_launches(st, 0, nullptr));
  if (far_used) LLIE_CUDA_OK(cudaStreamWaitEvent(st, ctx->sh_ev[0][0], 0));
  const uint64_t key = tile_key(ctx, d, n_dev, in[d], nullptr, nullptr, out[d], 0, g, H, W, p);
  return run_as_graph(ctx, ctx->g_tile[1], key, st, [&](cudaStream_t s2) -> int {
    int n_b = 0;
    for (auto& sp : pl->scales) n_b += sp.level;
    if (n_b) {
      LLIE_LAUNCH(ctx, s2, "decim4_x_kernel", decim4_x_kernel<<<dim3(ceil_div(wB, kDecXTile), g.c1_hi - g.c1_lo, 3), 128, 0, s2>>>(b.C, b.TB, hC, wC, wB, g.c1_lo, g.c1_hi));
      LLIE_LAUNCH(ctx, s2, "decim4_y_kernel", decim4_y_kernel<<<dim3(ceil_div(wB, 128), ceil_div(g.bb_hi - g.bb_lo, kDecQ), 3), 128, 0, s2>>>(b.TB, b.B, hC, hB, wB, g.bb_lo, g.bb_hi));
    }
    std::vector<int> b_scales;
    LLIE_TRY(scale_launches(s2, 1, &b_scales));
    if (!b_scales.empty())
      LLIE_TRY(launch_up_bc_y(ctx, s2, pl, b_scales, b.T2, t2_elems, b.SC, c_elems, hB, hC, wC, g.s_lo, g.s_hi, 3));
    const ImgView v = make_view(LLIE_NHWC, 3, g.R1 - g.R0, W);
    const bool vec = (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(in[d]) | reinterpret_cast<uintptr_t>(out[d])) % 16 == 0);
    return final_passes(ctx, s2, pl, p, in[d], out[d], v, vec, true, 1, g.R1 - g.R0, W, g.R0, g.R0, g.R1, hC, wC, b.SC, c_elems,
                        nullptr);
  });
}

// Reusable rendezvous of the per-device host threads.  The threads are dedicated (one per ctx) and the sections
// between rendezvous last tens of microseconds, so they spin (a futex wake costs more than a whole section) and
// only yield when a peer is late.
struct HostBarrier {
  std::atomic<int> waiting{0}, generation{0};
  int n;
  explicit HostBarrier(int n_) : n(n_) {}
  void wait() {
    const int gen = generation.load(std::memory_order_acquire);
    if (waiting.fetch_add(1, std::memory_order_acq_rel) + 1 == n) {
      waiting.store(0, std::memory_order_relaxed);
      generation.fetch_add(1, std::memory_order_release);
    } else {
      for (int spins = 0; generation.load(std::memory_order_acquire) == gen; ++spins)
        if (spins > 20000) std::this_thread::yield();
    }
  }
};

}  // namespace

int msrcr_tiled_run(llie_ctx* const* ctxs, int n_dev, const float* const* in, float* const* out, const int* row0,
                    const int* rows, int H, int W, const llie_msrcr_params& p) {
  LLIE_REQUIRE(p.n_scales <= 3, LLIE_E_ARG, "row-tiled MSRCR supports at most 3 scales (got %d)", p.n_scales);
  // plans (pure functions of the params) first, on the calling thread, so that errors surface here
  int r0_max = 0, rb_max = 0;
  bool any_b = false;
  for (int k = 0; k < p.n_scales; ++k) {
    ScalePlan sp;
    LLIE_TRY(get_scale_plan(p.sigma[k], p.truncate, &sp));
    if (sp.level) { rb_max = std::max(rb_max, sp.radius_low); any_b = true; }
    else r0_max = std::max(r0_max, sp.radius_low);
  }
  const int hC = level_c_size(H), wC = level_c_size(W), hB = level_b_size(H), wB = level_b_size(W);
  const int RW = 3 * W;
  std::vector<TileGeom> geo(n_dev);
  for (int d = 0; d < n_dev; ++d) geo[d] = tile_geometry(d, n_dev, row0, rows, H, r0_max, rb_max, any_b);
  std::vector<TileBufs> bufs(n_dev);
  std::vector<cudaEvent_t> ev_c(n_dev, nullptr);
  std::vector<int> rc_a(n_dev, LLIE_OK), rc(n_dev, LLIE_OK);
  std::vector<std::string> msg(n_dev);
  HostBarrier bar(n_dev);
  const bool percentile = p.finish == LLIE_FINISH_PERCENTILE;
  const size_t hb = msrcr_select_hist_bytes();

  // One PERSISTENT host thread per device (the ctx's HostWorker; the caller's thread drives shard 0): the ~25
  // launches and copies of a device must not serialise behind the other devices', and spawning threads per call
  // cost more than the whole 8-GPU step.  Two spin rendezvous per call: "set up" and "C-ready events recorded".
  auto worker = [&](int d) {
    llie_ctx* ctx = ctxs[d];
    auto setup = [&]() -> int {
      LLIE_CUDA_OK(cudaSetDevice(ctx->device));
      LLIE_TRY(ensure_plan(ctx, p, ctx->stream));
      LLIE_TRY(sharded_ensure_streams(ctx));  // the copy stream + its event (shared with the batch sharder)
      TileBufs& b = bufs[d];
      const size_t c_bytes = (size_t)3 * hC * wC * 4, b_bytes = (size_t)3 * hB * wB * 4;
      void* q;
      LLIE_TRY(ctx->workspace("tile.V", (size_t)hC * RW * 4, &q)); b.V = (float*)q;
      LLIE_TRY(ctx->workspace("tile.C", c_bytes, &q)); b.C = (float*)q;
      LLIE_TRY(ctx->workspace("tile.Ct", c_bytes, &q)); b.Ct = (float*)q;
      LLIE_TRY(ctx->workspace("tile.SC", c_bytes * 3, &q)); b.SC = (float*)q;
      LLIE_TRY(ctx->workspace("tile.TB", (size_t)3 * hC * wB * 4, &q)); b.TB = (float*)q;
      LLIE_TRY(ctx->workspace("tile.B", b_bytes, &q)); b.B = (float*)q;
      LLIE_TRY(ctx->workspace("tile.Bt", b_bytes, &q)); b.Bt = (float*)q;
      LLIE_TRY(ctx->workspace("tile.SB", b_bytes, &q)); b.SB = (float*)q;
      LLIE_TRY(ctx->workspace("tile.T2", (size_t)3 * 3 * hB * wC * 4, &q)); b.T2 = (float*)q;  // one per scale (<= 3)
      LLIE_TRY(ctx->workspace("tile.band_top", (size_t)kBandTopRows * RW * 4, &q)); b.band_top = (float*)q;
      LLIE_TRY(ctx->workspace("tile.band_bot", (size_t)kBandBotRows * RW * 4, &q)); b.band_bot = (float*)q;
      if (p.finish != LLIE_FINISH_NONE) {
        LLIE_TRY(ctx->workspace("tile.finish.partials", std::max(msrcr_finish_partials_bytes(), hb), &b.fin_part));
        LLIE_TRY(ctx->workspace("tile.finish.gather", std::max(msrcr_finish_partials_bytes(), 3 * hb) * n_dev, &b.fin_gather));
        LLIE_TRY(ctx->workspace("tile.finish.lo_inv", 3 * sizeof(float2), &b.fin_li));
        LLIE_TRY(ctx->workspace("tile.finish.state", msrcr_select_state_bytes(), &b.fin_state));
      }
      if (!ctx->tile_event) {
        LLIE_CUDA_OK(cudaEventCreateWithFlags(&ctx->tile_event, cudaEventDisableTiming));
        LLIE_CUDA_OK(cudaEventCreate(&ctx->tile_t0));
        LLIE_CUDA_OK(cudaEventCreate(&ctx->tile_t1));
      }
      ev_c[d] = ctx->tile_event;
      ctx->tile_ms = -1.f;
      return LLIE_OK;
    };
    auto phase_a = [&]() -> int {
      LLIE_CUDA_OK(cudaEventRecord(ctx->tile_t0, ctx->stream));
      LLIE_TRY(tiled_phase_a(ctx, d, n_dev, ctxs, in, rows, geo[d], bufs[d], H, W, p));
      LLIE_CUDA_OK(cudaEventRecord(ev_c[d], ctx->stream));
      return LLIE_OK;
    };
    auto fail = [&](int code) {
      if (rc[d] == LLIE_OK && code != LLIE_OK) {
        rc[d] = code;
        msg[d] = last_error_text();
      }
    };
    rc_a[d] = setup();
    bar.wait();  // common start; no device allocates (first call, or a larger image) while another one captures a graph
    if (rc_a[d] == LLIE_OK) rc_a[d] = phase_a();
    fail(rc_a[d]);
    bar.wait();  // every device's C-ready event is recorded (or its phase failed)
    bool ok = true;  // rc_a[] is final after the rendezvous; rc[] is only read by its owner until the end
    for (int e = 0; e < n_dev; ++e) ok = ok && rc_a[e] == LLIE_OK;
    if (ok) fail(tiled_phase_b(ctx, d, n_dev, ctxs, in, out, geo, bufs, ev_c, H, W, p));
    if (p.finish != LLIE_FINISH_NONE) {
      // Global-statistics finisher across the shards ("one tiny all-reduce"): every device
      // reduces its own band to partials and pushes them into every device's gather buffer (an all-gather of
      // 16-48 KB blocks by peer copies); each device then folds all partials in the same order -- identical
      // (lo, 1/(hi-lo)) everywhere -- and stretches its band in place.  The percentile finisher does this three
      // times (one radix-select pass each: 11 + 11 + 10 bits) on per-band histograms.
      // Every round: [reduce, push, record] -> rendezvous -> [wait, fold]; a second rendezvous separates "everybody
      // has enqueued its waits" from the next round's re-recording of the events.
      const int band_rows = geo[d].R1 - geo[d].R0;
      const int rounds = percentile ? 3 : 1;
      std::vector<int> okv(n_dev);
      for (int round = 0; round < rounds; ++round) {
        const size_t pb = percentile ? hb : msrcr_finish_partials_bytes();
        auto stage1 = [&]() -> int {
          if (percentile) LLIE_TRY(msrcr_select_hist(ctx, ctx->stream, out[d], band_rows, W, round, bufs[d].fin_state, bufs[d].fin_part));
          else LLIE_TRY(msrcr_finish_stats(ctx, ctx->stream, out[d], 1, band_rows, W, LLIE_NHWC, bufs[d].fin_part));
          for (int e = 0; e < n_dev; ++e)
            LLIE_TRY(copy_dev((char*)bufs[e].fin_gather + ((size_t)round * n_dev + d) * pb, ctxs[e]->device, bufs[d].fin_part,
                              ctx->device, pb, ctx->stream));
          LLIE_CUDA_OK(cudaEventRecord(ev_c[d], ctx->stream));
          return LLIE_OK;
        };
        auto stage2 = [&]() -> int {
          for (int e = 0; e < n_dev; ++e)
            if (e != d) LLIE_CUDA_OK(cudaStreamWaitEvent(ctx->stream, ev_c[e], 0));
          const char* gather = (const char*)bufs[d].fin_gather + (size_t)round * n_dev * pb;
          if (percentile) {
            LLIE_TRY(msrcr_select_scan(ctx, ctx->stream, round, gather, n_dev, bufs[d].fin_state, bufs[d].fin_li, (long long)H * W, p.finish_k));
            if (round == 2) LLIE_TRY(msrcr_stretch(ctx, ctx->stream, out[d], band_rows, W, bufs[d].fin_li));
            return LLIE_OK;
          }
          return msrcr_finish_apply(ctx, ctx->stream, out[d], 1, band_rows, W, LLIE_NHWC, gather,
                                    n_dev * (int)(pb / msrcr_finish_partial_size()), (double)H * W, p, bufs[d].fin_li);
        };
        bar.wait();  // every device has enqueued its waits on the events before anyone records them again
        if (ok && rc[d] == LLIE_OK) fail(stage1());
        rc_a[d] = rc[d];
        bar.wait();  // all pushes are enqueued and their events recorded
        for (int e = 0; e < n_dev; ++e) ok = ok && rc_a[e] == LLIE_OK;
        if (ok) fail(stage2());
      }
    }
    if (ok && rc[d] == LLIE_OK) cudaEventRecord(ctx->tile_t1, ctx->stream);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess && ctx->sh_h2d) e = cudaStreamSynchronize(ctx->sh_h2d);
    if (e == cudaSuccess && ok && rc[d] == LLIE_OK) cudaEventElapsedTime(&ctx->tile_ms, ctx->tile_t0, ctx->tile_t1);
    if (e != cudaSuccess && rc[d] == LLIE_OK) {
      rc[d] = LLIE_E_CUDA;
      msg[d] = std::string("tiled MSRCR: device ") + std::to_string(ctx->device) + ": " + cudaGetErrorString(e);
    }
  };
  int prev_dev = 0;
  cudaGetDevice(&prev_dev);
  for (int d = 1; d < n_dev; ++d) ctx_worker(ctxs[d])->post([&worker, d] { worker(d); });
  worker(0);
  // nobody returns (and no C buffer is reused) before every device has finished copying from its neighbours
  for (int d = 1; d < n_dev; ++d) ctxs[d]->worker->wait();
  cudaSetDevice(prev_dev);
  for (int d = 0; d < n_dev; ++d)
    if (rc[d] != LLIE_OK) {
      set_error("%s", msg[d].c_str());
      return rc[d];
    }
  return LLIE_OK;
}

}  // namespace llie
