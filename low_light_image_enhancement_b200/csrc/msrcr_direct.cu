// MSRCR, exact path (LLIE_BLUR_DIRECT): the truncated Gaussian FIR of SURVEY.md 9.1 evaluated
// tap by tap in fp32.  O(sigma) work per pixel -- this is the GPU-side ground truth that the
// O(1) pyramid path (msrcr_pyramid.cu) is checked against, and the path for callers who need
// the oracle's hard 4-sigma truncation reproduced exactly.
//
// Hot-path units: a1 Gaussian surround, a2 log-ratio, a3 colour restoration,
// a4 gain/offset.  No reference file exists to cite (/root/reference/README.md:1-2).
#include "fir.cuh"

#include <cmath>

namespace llie {
namespace {

struct EpiParams {
  int n_scales;
  float weight[LLIE_MAX_SCALES];
  float alpha, beta, gain, offset, in_scale, in_bias;
};

// Per pixel: R_c = sum_s w_s (ln I_c - ln S_{s,c}); C_c = beta (ln(alpha I_c) - ln sum_c I_c);
// out = clip((G C R + b) / 255).  surround[s] has the image's own layout.
__global__ void msrcr_epilogue_kernel(const float* __restrict__ x, const float* __restrict__ surround,
                                      long long scale_stride, float* __restrict__ out, long long n_px_total,
                                      long long px_per_img, ImgView v, int W, EpiParams p) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n_px_total) return;
  long long img = i / px_per_img, pix = i - img * px_per_img;
  long long yy = pix / W, xx = pix - yy * W;
  long long base = img * v.sn + yy * v.sy + xx * v.sx;
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
    float Cc = p.beta * (lalpha + lI[c] - lsum);
    float o = fmaf(p.gain * Cc, R, p.offset) * (1.0f / 255.0f);
    out[base + c * v.sc] = fminf(fmaxf(o, 0.f), 1.f);
  }
}

}  // namespace

int msrcr_direct_run(llie_ctx* ctx, const float* in, float* out, int n, int h, int w, int layout,
                     const llie_msrcr_params& p, cudaStream_t st) {
  const int C = (layout == LLIE_NHWC) ? 3 : 1;
  const long long img_elems = (long long)h * w * 3;
  // taps for all scales, computed in double like the oracle and rounded once to fp32
  std::vector<float> taps;
  std::vector<int> radius(p.n_scales), tap_off(p.n_scales);
  int rmax = 0;
  for (int s = 0; s < p.n_scales; ++s) {
    int r = (int)std::ceil((double)p.truncate * (double)p.sigma[s]);
    radius[s] = r;
    rmax = r > rmax ? r : rmax;
    std::vector<double> k(2 * r + 1);
    double sum = 0;
    for (int t = -r; t <= r; ++t) sum += (k[t + r] = std::exp(-(double)t * t / (2.0 * p.sigma[s] * p.sigma[s])));
    tap_off[s] = (int)taps.size();
    for (double v : k) taps.push_back((float)(v / sum));
    taps.resize((taps.size() + 3) & ~size_t(3));
  }
  size_t smem = (size_t)(128 + 2 * rmax + 3) * C * sizeof(float);
  LLIE_REQUIRE(smem <= 200 * 1024, LLIE_E_ARG, "direct blur: radius %d too large for shared-memory staging", rmax);
  LLIE_CUDA_OK(cudaFuncSetAttribute(fir_x_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));

  // frames per pass bounded so that (1 + n_scales) temporaries stay under ~1 GiB
  long long chunk = (1ll << 30) / ((1 + p.n_scales) * img_elems * 4);
  chunk = chunk < 1 ? 1 : (chunk > n ? n : chunk);
  void *d_taps, *d_tmp, *d_sur;
  LLIE_TRY(ctx->workspace("msrcr.direct.taps", taps.size() * sizeof(float), &d_taps));
  LLIE_TRY(ctx->workspace("msrcr.direct.tmp", (size_t)chunk * img_elems * 4, &d_tmp));
  LLIE_TRY(ctx->workspace("msrcr.direct.surround", (size_t)chunk * img_elems * 4 * p.n_scales, &d_sur));
  LLIE_CUDA_OK(cudaMemcpyAsync(d_taps, taps.data(), taps.size() * sizeof(float), cudaMemcpyHostToDevice, st));
  // the pageable host vector must outlive the copy
  LLIE_CUDA_OK(cudaStreamSynchronize(st));

  EpiParams ep;
  ep.n_scales = p.n_scales;
  for (int s = 0; s < p.n_scales; ++s) ep.weight[s] = p.weight[s];
  ep.alpha = p.alpha; ep.beta = p.beta; ep.gain = p.gain; ep.offset = p.offset;
  ep.in_scale = p.in_scale; ep.in_bias = p.in_bias;
  const ImgView v = make_view(layout, 3, h, w);

  for (long long n0 = 0; n0 < n; n0 += chunk) {
    const int nb = (int)((n - n0 < chunk) ? (n - n0) : chunk);
    const float* src = in + n0 * img_elems;
    const long long rows = (long long)nb * h * (3 / C);     // NHWC: nb*h rows of 3W; NCHW: nb*3*h rows of W
    const int planes = nb * (3 / C), RW = w * C;
    for (int s = 0; s < p.n_scales; ++s) {
      const float* k = (const float*)d_taps + tap_off[s];
      float* sur = (float*)d_sur + (size_t)s * chunk * img_elems;
      dim3 gx((unsigned)rows, (w + 127) / 128);
      size_t sm = (size_t)(128 + 2 * radius[s] + 3) * C * sizeof(float);
      LLIE_LAUNCH(ctx, st, "fir_x_kernel", fir_x_kernel<true><<<gx, 128, sm, st>>>(src, (float*)d_tmp, k, radius[s], w, C, p.in_scale, p.in_bias, (int)rows, 0, 0));
      dim3 gy((RW + 127) / 128, (h + 3) / 4, planes);
      LLIE_LAUNCH(ctx, st, "fir_y_kernel", fir_y_kernel<<<gy, 128, 0, st>>>((const float*)d_tmp, sur, k, radius[s], h, RW, 0, h, 0, h, 0, h));
    }
    long long npx = (long long)nb * h * w;
    LLIE_LAUNCH(ctx, st, "msrcr_epilogue_kernel", msrcr_epilogue_kernel<<<(unsigned)((npx + 255) / 256), 256, 0, st>>>(
        src, (const float*)d_sur, chunk * img_elems, out + n0 * img_elems, npx, (long long)h * w, v, w, ep));
  }
  return LLIE_OK;
}

}  // namespace llie
