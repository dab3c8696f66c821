// Zero-DCE, fp32 path (config 1: "random-init fp32", tolerance 1e-4): DCE-Net as seven direct
// 3x3 convolutions on the CUDA cores with fp32 FMA, plus the fused tanh/curve pass.
//
// Hot-path units: a5 DCE-Net, a6 curve iterations; network spec SURVEY.md 9.2.
// No reference file exists to cite (/root/reference/README.md:1-2).
//
// Kernel shape: a CTA owns a 64x4 pixel tile and all output channels; each thread keeps a
// 4 pixel x 8 channel register tile (32 accumulators) and walks the input channels in chunks of
// 8 that are staged in shared memory with cp.async double buffering.  Per (cin, dy) a thread
// issues 1.5 input LDS.128 + 6 broadcast weight LDS.128 for 96 FMAs, so the FMA pipe is the limit.
#include "dce_weights.cuh"

#include <cmath>
#include <cstring>

namespace llie {

void dce_weights_free(DceWeights* w) {
  if (!w) return;
  cudaFree(w->d_f32);
  cudaFree(w->d_bf16);
  delete w;
}

int dce_weights_upload(llie_ctx* ctx, const float* host_packed, size_t count) {
  LLIE_REQUIRE(host_packed != nullptr, LLIE_E_ARG, "dce_set_weights: null blob");
  LLIE_REQUIRE(count == kNumParams, LLIE_E_SHAPE, "dce_set_weights: expected %zu floats, got %zu", kNumParams, count);
  if (ctx->dce) { dce_weights_free(ctx->dce); ctx->dce = nullptr; }
  DceWeights* dw = new DceWeights;
  dw->host_raw.assign(host_packed, host_packed + count);
  std::vector<float> pk;
  size_t src = 0;
  for (int l = 0; l < 7; ++l) {
    const int co = kCout[l], ci = kCin[l], cip = (ci + 7) & ~7;
    dw->w_off[l] = pk.size();
    pk.resize(pk.size() + (size_t)cip * 9 * 32, 0.f);
    float* dst = pk.data() + dw->w_off[l];
    for (int o = 0; o < co; ++o)
      for (int i = 0; i < ci; ++i)
        for (int t = 0; t < 9; ++t) dst[((size_t)i * 9 + t) * 32 + o] = host_packed[src + ((size_t)o * ci + i) * 9 + t];
    src += (size_t)co * ci * 9;
    dw->b_off[l] = pk.size();
    pk.resize(pk.size() + 32, 0.f);
    for (int o = 0; o < co; ++o) pk[dw->b_off[l] + o] = host_packed[src + o];
    src += co;
  }
  LLIE_CUDA_OK(cudaMalloc(&dw->d_f32, pk.size() * sizeof(float)));
  LLIE_CUDA_OK(cudaMemcpy(dw->d_f32, pk.data(), pk.size() * sizeof(float), cudaMemcpyHostToDevice));
  ctx->dce = dw;
  return dce_pack_bf16(ctx, dw);
}

namespace {

constexpr int TW = 64, TH = 4, CK = 8;
constexpr int SROW = 68;                       // 66 used columns, padded to a float4 multiple
constexpr int IN_STAGE = CK * (TH + 2) * SROW; // floats
constexpr int W_STAGE = CK * 9 * 32;           // floats

__device__ __forceinline__ void cp_async4_zfill(float* smem_dst, const float* gsrc, bool valid) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  int sz = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(gsrc), "r"(sz));
}
__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// NG = number of 8-wide output-channel groups (4 -> 32 outputs, 3 -> 24 outputs)
// ACT: 0 = ReLU, 1 = tanh.  Output is planar fp32 [n][NG*8][H][W].
template <int NG, int ACT>
__global__ void __launch_bounds__(NG * 64)
dce_conv3x3_f32_kernel(const float* __restrict__ in_a, ImgView va, int ca, const float* __restrict__ in_b, ImgView vb,
                       int cb, const float* __restrict__ wpk, const float* __restrict__ bias, float* __restrict__ out,
                       int H, int W) {
  extern __shared__ __align__(16) float smem[];
  float* s_in = smem;                   // [2][CK][TH+2][SROW]
  float* s_w = smem + 2 * IN_STAGE;     // [2][CK][9][32]
  const int tid = threadIdx.x;
  const int cg = tid >> 6, pt = tid & 63, pi = pt & 15, ry = pt >> 4;
  const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
  const long long img = blockIdx.z;
  const int ca_pad = (ca + CK - 1) / CK * CK;
  const int n_chunks = (ca_pad + cb) / CK;

  auto load_chunk = [&](int chunk, int stage) {
    const int c0 = chunk * CK;
    const bool from_b = c0 >= ca_pad;
    const float* base = from_b ? in_b + img * vb.sn : in_a + img * va.sn;
    const ImgView v = from_b ? vb : va;
    const int cbase = from_b ? c0 - ca_pad : c0;
    const int climit = from_b ? cb : ca;
    float* di = s_in + stage * IN_STAGE;
    for (int i = tid; i < CK * (TH + 2) * 66; i += NG * 64) {
      int c = i / ((TH + 2) * 66), rem = i - c * (TH + 2) * 66;
      int rr = rem / 66, xx = rem - rr * 66;
      int gy = y0 - 1 + rr, gx = x0 - 1 + xx, gc = cbase + c;
      bool ok = (gy >= 0) && (gy < H) && (gx >= 0) && (gx < W) && (gc < climit);
      const float* g = ok ? base + gc * v.sc + gy * v.sy + gx * v.sx : base;
      cp_async4_zfill(di + (c * (TH + 2) + rr) * SROW + xx, g, ok);
    }
    const float* wsrc = wpk + (size_t)c0 * 9 * 32;
    float* dw = s_w + stage * W_STAGE;
    for (int i = tid; i < W_STAGE / 4; i += NG * 64) cp_async16(dw + i * 4, wsrc + i * 4);
  };

  float acc[4][8];
#pragma unroll
  for (int p = 0; p < 4; ++p)
#pragma unroll
    for (int o = 0; o < 8; ++o) acc[p][o] = 0.f;

  load_chunk(0, 0);
  cp_async_commit();
  for (int ch = 0; ch < n_chunks; ++ch) {
    if (ch + 1 < n_chunks) load_chunk(ch + 1, (ch + 1) & 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    const float* si = s_in + (ch & 1) * IN_STAGE;
    const float* sw = s_w + (ch & 1) * W_STAGE + cg * 8;
#pragma unroll 2
    for (int c = 0; c < CK; ++c) {
#pragma unroll
      for (int dy = 0; dy < 3; ++dy) {
        const float* rowp = si + (c * (TH + 2) + ry + dy) * SROW + pi * 4;
        float4 a = *reinterpret_cast<const float4*>(rowp);
        float2 b = *reinterpret_cast<const float2*>(rowp + 4);
        float iv[6] = {a.x, a.y, a.z, a.w, b.x, b.y};
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
          const float* wp = sw + (c * 9 + dy * 3 + dx) * 32;
          float4 w0 = *reinterpret_cast<const float4*>(wp);
          float4 w1 = *reinterpret_cast<const float4*>(wp + 4);
          float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
          for (int p = 0; p < 4; ++p)
#pragma unroll
            for (int o = 0; o < 8; ++o) acc[p][o] = fmaf(iv[p + dx], wv[o], acc[p][o]);
        }
      }
    }
    __syncthreads();
  }

  const int y = y0 + ry, x = x0 + pi * 4;
  if (y >= H || x >= W) return;
  const bool vec = ((W & 3) == 0);
#pragma unroll
  for (int o = 0; o < 8; ++o) {
    const int co = cg * 8 + o;
    const float bv = bias[co];
    float r[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      float t = acc[p][o] + bv;
      r[p] = ACT == 0 ? fmaxf(t, 0.f) : tanhf(t);
    }
    float* dst = out + ((img * (NG * 8) + co) * H + y) * (long long)W + x;
    if (vec) {
      *reinterpret_cast<float4*>(dst) = make_float4(r[0], r[1], r[2], r[3]);
    } else {
#pragma unroll
      for (int p = 0; p < 4; ++p)
        if (x + p < W) dst[p] = r[p];
    }
  }
}

// north_star kernel (3): LE iterations y <- y + s*A_n*y*(1-y) on the 24 curve maps, one pass.
// maps are read through `vm` (24 channels); optionally re-emitted in the caller's layout.
__global__ void dce_curves_kernel(const float* __restrict__ x, ImgView vx, const float* __restrict__ maps, ImgView vm,
                                  float* __restrict__ out, float* __restrict__ maps_out, ImgView vmo,
                                  long long n_px_total, long long px_per_img, int W, int iters, float sign) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n_px_total) return;
  long long img = i / px_per_img, pix = i - img * px_per_img;
  long long yy = pix / W, xx = pix - yy * W;
  const long long bx = img * vx.sn + yy * vx.sy + xx * vx.sx;
  const long long bm = img * vm.sn + yy * vm.sy + xx * vm.sx;
  float a[24];
#pragma unroll
  for (int c = 0; c < 24; ++c) a[c] = maps[bm + c * vm.sc];
  float v[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) v[c] = x[bx + c * vx.sc];
#pragma unroll
  for (int n = 0; n < 8; ++n) {
    if (n < iters) {
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] = fmaf(sign * a[3 * n + c] * v[c], 1.0f - v[c], v[c]);
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) out[bx + c * vx.sc] = v[c];
  if (maps_out) {
    const long long bo = img * vmo.sn + yy * vmo.sy + xx * vmo.sx;
#pragma unroll
    for (int c = 0; c < 24; ++c) maps_out[bo + c * vmo.sc] = a[c];
  }
}

template <int NG, int ACT>
int launch_conv(llie_ctx* ctx, cudaStream_t st, const float* a, ImgView va, int ca, const float* b, ImgView vb, int cb,
                const float* wpk, const float* bias, float* out, int n, int h, int w) {
  constexpr size_t smem = (2 * IN_STAGE + 2 * W_STAGE) * sizeof(float);
  LLIE_CUDA_OK(cudaFuncSetAttribute(dce_conv3x3_f32_kernel<NG, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)smem));
  dim3 grid((w + TW - 1) / TW, (h + TH - 1) / TH, n);
  LLIE_LAUNCH(ctx, st, "dce_conv3x3_f32_kernel",
              (dce_conv3x3_f32_kernel<NG, ACT><<<grid, NG * 64, smem, st>>>(a, va, ca, b, vb, cb, wpk, bias, out, h, w)));
  return LLIE_OK;
}

}  // namespace

int dce_curves_run(llie_ctx* ctx, const float* in, const float* maps, ImgView vm, float* out, float* maps_out, int n,
                   int h, int w, int layout, const llie_dce_params& p, cudaStream_t st) {
  const ImgView vx = make_view(layout, 3, h, w);
  const ImgView vmo = make_view(layout, 24, h, w);
  long long npx = (long long)n * h * w;
  LLIE_LAUNCH(ctx, st, "dce_curves_kernel",
              dce_curves_kernel<<<(unsigned)((npx + 127) / 128), 128, 0, st>>>(
                  in, vx, maps, vm, out, maps_out, vmo, npx, (long long)h * w, w, p.iters,
                  p.curve_sign >= 0 ? 1.f : -1.f));
  return LLIE_OK;
}

int dce_forward_f32(llie_ctx* ctx, const float* in, float* out, float* maps, int n, int h, int w, int layout,
                    const llie_dce_params& p, cudaStream_t st) {
  DceWeights* dw = ctx->dce;
  LLIE_REQUIRE(dw != nullptr, LLIE_E_STATE, "dce_forward: call llie_dce_set_weights first");
  const long long hw = (long long)h * w;
  // frames per pass: six 32-channel fp32 activations + the 24 maps, kept under ~2 GiB
  long long chunk = (2ll << 30) / ((6 * 32 + 24) * hw * 4);
  chunk = chunk < 1 ? 1 : (chunk > n ? n : chunk);
  void* ws;
  LLIE_TRY(ctx->workspace("dce.f32.act", (size_t)chunk * (6 * 32 + 24) * hw * 4, &ws));
  float* act[6];
  for (int i = 0; i < 6; ++i) act[i] = (float*)ws + (size_t)i * chunk * 32 * hw;
  float* A = (float*)ws + (size_t)6 * chunk * 32 * hw;
  const ImgView vin = make_view(layout, 3, h, w);
  const ImgView v32 = make_view(LLIE_NCHW, 32, h, w);
  const ImgView v24 = make_view(LLIE_NCHW, 24, h, w);
  const float* W = dw->d_f32;
  for (long long n0 = 0; n0 < n; n0 += chunk) {
    const int nb = (int)((n - n0 < chunk) ? (n - n0) : chunk);
    const float* x = in + n0 * hw * 3;
#define LLIE_CONV(NG, ACT, l, a, va, ca, b, cb, o) \
  LLIE_TRY((launch_conv<NG, ACT>(ctx, st, a, va, ca, b, v32, cb, W + dw->w_off[l], W + dw->b_off[l], o, nb, h, w)))
    LLIE_CONV(4, 0, 0, x, vin, 3, nullptr, 0, act[0]);
    LLIE_CONV(4, 0, 1, act[0], v32, 32, nullptr, 0, act[1]);
    LLIE_CONV(4, 0, 2, act[1], v32, 32, nullptr, 0, act[2]);
    LLIE_CONV(4, 0, 3, act[2], v32, 32, nullptr, 0, act[3]);
    LLIE_CONV(4, 0, 4, act[2], v32, 32, act[3], 32, act[4]);
    LLIE_CONV(4, 0, 5, act[1], v32, 32, act[4], 32, act[5]);
    LLIE_CONV(3, 1, 6, act[0], v32, 32, act[5], 32, A);
#undef LLIE_CONV
    LLIE_TRY(dce_curves_run(ctx, x, A, v24, out + n0 * hw * 3, maps ? maps + n0 * hw * 24 : nullptr, nb, h, w, layout,
                            p, st));
  }
  return LLIE_OK;
}

}  // namespace llie
