// Device kernels of the MSRCR fast path (LLIE_BLUR_PYRAMID): a two-level decimated Gaussian cascade
// whose cost per pixel does not depend on sigma (north_star kernel (1)).
// No reference file exists to cite (/root/reference/README.md:1-2).
//
// Levels (1-D view; the 2-D operators are separable):
//   C (stride 4):  C[j] = sum_t g[t] I[clamp(4(j-PC)+t-16)],  g = Gaussian sigma 4, radius 16
//   B (stride 16): B[j] = sum_t g[t] C[clamp(4(j-PB)+PC+t-16)]   (same g, in C samples)
// A scale with sigma < 40 is blurred on C, a larger one on B and then brought to C resolution by
// a Gaussian-weighted x4 upsample (unit sigma in B samples, 9 taps; its variance is part of the
// budget).  The last x4 step C -> pixels is Lagrange interpolation (6 taps for C-level scales,
// 4 for B-level ones) fused with the log / ratio / colour-restoration / gain-offset epilogue.
// Edge clamping is exact: each level stores PC / PB samples beyond the border, which reaches the
// region where the clamped, prefiltered signal is constant.
#pragma once
#include "llie_common.cuh"

namespace llie {
namespace {

constexpr int kPC = 4;   // C samples stored before pixel 0
constexpr int kPB = 5;   // B samples stored before pixel 0
constexpr int kRD = 16;  // radius of the decimation prefilter (33 taps)
constexpr int kTU = 4;   // half-width of the B -> C Gaussian upsampler (9 taps)

__host__ __device__ inline int level_c_size(int n) { return kPC + (n + 18) / 4 + 1; }
__host__ __device__ inline int level_b_size(int n) { return kPB + (n + 94) / 16 + 1; }

// Gaussian sigma = 4, radius 16, sum-normalised; g[t] = g[32 - t], first 17 taps listed.
// Compile-time constants so that fully unrolled filters use immediate-operand FFMAs.
__device__ constexpr float kGD[17] = {
    3.345874043e-05f, 8.815204000e-05f, 2.181783944e-04f, 5.072799653e-04f, 1.108001311e-03f, 2.273471076e-03f,
    4.382230196e-03f, 7.935193759e-03f, 1.349821928e-02f, 2.157009300e-02f, 3.238054475e-02f, 4.566388724e-02f,
    6.049482184e-02f, 7.528702219e-02f, 8.801944640e-02f, 9.667045004e-02f, 9.973909953e-02f};
__device__ __forceinline__ constexpr float gd(int t) { return kGD[t <= 16 ? t : 32 - t]; }

// Lagrange interpolation weights at fraction ph/4 (ph = 1..3); ph = 0 is the sample itself.
// 6 taps at offsets -2..3, 4 taps at offsets -1..2.
__device__ constexpr float kL6[3][6] = {
    {9.3994140625e-03f, -8.4594726563e-02f, 8.4594726562e-01f, 2.8198242188e-01f, -6.0424804688e-02f, 7.6904296875e-03f},
    {1.1718750000e-02f, -9.7656250000e-02f, 5.8593750000e-01f, 5.8593750000e-01f, -9.7656250000e-02f, 1.1718750000e-02f},
    {7.6904296875e-03f, -6.0424804688e-02f, 2.8198242188e-01f, 8.4594726562e-01f, -8.4594726562e-02f, 9.3994140625e-03f}};
__device__ constexpr float kL4[3][4] = {{-5.4687500000e-02f, 8.2031250000e-01f, 2.7343750000e-01f, -3.9062500000e-02f},
                                        {-6.2500000000e-02f, 5.6250000000e-01f, 5.6250000000e-01f, -6.2500000000e-02f},
                                        {-3.9062500000e-02f, 2.7343750000e-01f, 8.2031250000e-01f, -5.4687500000e-02f}};

// ---- K1: vertical prefilter + decimate by 4, full resolution -> V ------------------------------
// The image is `planes` planes of rows of RW floats (NHWC: one plane per image, RW = 3W, channels
// interleaved; NCHW: three planes per image, RW = W).  V[plane][jy][e] = sum_t g[t] x[clamp(4(jy-PC)
// + t - 16, 0, H-1)][e] on the raw input (the affine map I = a x + b commutes with a sum-1 filter
// and is applied by K1b).  A thread owns VEC consecutive floats of the row and walks down the raw
// rows once, scattering each loaded value into the 8-9 output rows whose windows contain it; the
// 9 live accumulators rotate through a 9-group (36-row) unrolled body so every index is static.
// `x` points at image row in_row0 and holds in_rows rows (row-tiled shards); rows outside [0, H) clamp.
// TIn = float, or uint8_t (8-bit images; raw 0..255 values are filtered as they are and
// the 1/255 is folded into K1b's affine map).
template <int VEC, typename TIn>
__global__ void __launch_bounds__(128)
down4_rows_kernel(const TIn* __restrict__ x, long long plane_stride, int RW, int H, int in_row0, int in_rows,
                  float* __restrict__ V, int hC, int v_row_begin, int v_row_end, int rows_per_strip) {
  const int e0 = (blockIdx.x * 128 + threadIdx.x) * VEC;
  if (e0 >= RW) return;
  const int j0 = v_row_begin + blockIdx.y * rows_per_strip;
  const int j1 = min(j0 + rows_per_strip, v_row_end);
  if (j0 >= j1) return;
  const long long plane = blockIdx.z;
  const TIn* src = x + plane * plane_stride + e0;
  float* dst = V + plane * (long long)hC * RW + e0;
  const int ngroups = (j1 - j0) + 8;
  const int ybase = 4 * (j0 - kPC - 4);  // raw row of group 0, phase 0

  float acc[9][VEC];
#pragma unroll
  for (int s = 0; s < 9; ++s)
#pragma unroll
    for (int i = 0; i < VEC; ++i) acc[s][i] = 0.f;

  float cur[4][VEC], nxt[4][VEC];
  auto load_group = [&](float (&buf)[4][VEC], int g) {
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      // clamp to the image, then to the rows this pointer holds (the latter only ever bites on
      // taps of output rows outside [j0, j1), which are never stored)
      const int y = min(max(min(max(ybase + 4 * g + p, 0), H - 1) - in_row0, 0), in_rows - 1);
      const TIn* q = src + (long long)y * RW;
      if constexpr (VEC == 4 && sizeof(TIn) == 1) {
        const uchar4 t = __ldg(reinterpret_cast<const uchar4*>(q));
        buf[p][0] = (float)t.x; buf[p][1] = (float)t.y; buf[p][2] = (float)t.z; buf[p][3] = (float)t.w;
      } else if constexpr (VEC == 4 && sizeof(TIn) == 2) {
        const ushort4 t = __ldg(reinterpret_cast<const ushort4*>(q));
        buf[p][0] = (float)t.x; buf[p][1] = (float)t.y; buf[p][2] = (float)t.z; buf[p][3] = (float)t.w;
      } else if constexpr (VEC == 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(q));
        buf[p][0] = t.x; buf[p][1] = t.y; buf[p][2] = t.z; buf[p][3] = t.w;
      } else {
#pragma unroll
        for (int i = 0; i < VEC; ++i) buf[p][i] = (float)__ldg(q + i);
      }
    }
  };
  load_group(cur, 0);
  for (int gbase = 0; gbase < ngroups; gbase += 9) {
#pragma unroll
    for (int gg = 0; gg < 9; ++gg) {
      const int g = gbase + gg;
      if (g < ngroups) {
        if (g + 1 < ngroups) load_group(nxt, g + 1);
        // phase 0: taps 4q, q = 0..8, into slot (gg - q) mod 9
#pragma unroll
        for (int q = 0; q <= 8; ++q)
#pragma unroll
          for (int i = 0; i < VEC; ++i) acc[(gg - q + 9) % 9][i] = fmaf(gd(4 * q), cur[0][i], acc[(gg - q + 9) % 9][i]);
        {  // the row that just received its last tap: jy = j0 + g - 8, slot (gg + 1) mod 9
          const int jy = j0 + g - 8;
          if (g >= 8) {
            float* o = dst + (long long)jy * RW;
            if constexpr (VEC == 4) {
              *reinterpret_cast<float4*>(o) = make_float4(acc[(gg + 1) % 9][0], acc[(gg + 1) % 9][1],
                                                          acc[(gg + 1) % 9][2], acc[(gg + 1) % 9][3]);
            } else {
#pragma unroll
              for (int i = 0; i < VEC; ++i) o[i] = acc[(gg + 1) % 9][i];
            }
          }
#pragma unroll
          for (int i = 0; i < VEC; ++i) acc[(gg + 1) % 9][i] = 0.f;
        }
#pragma unroll
        for (int p = 1; p < 4; ++p)
#pragma unroll
          for (int q = 0; q <= 7; ++q)
#pragma unroll
            for (int i = 0; i < VEC; ++i)
              acc[(gg - q + 9) % 9][i] = fmaf(gd(4 * q + p), cur[p][i], acc[(gg - q + 9) % 9][i]);
#pragma unroll
        for (int p = 0; p < 4; ++p)
#pragma unroll
          for (int i = 0; i < VEC; ++i) cur[p][i] = nxt[p][i];
      }
    }
  }
}

// ---- K1b: horizontal prefilter + decimate by 4 on V -> planar level C, with I = a x + b ----------
// V[plane][jy][px*CS + c] -> C[(plane*CS + c)][jy][jx] = a * sum_t g[t] V[..][clamp(4(jx-PC)+t-16)] + b.
// A CTA stages one row segment (edge-clamped) in shared memory; thread jx reads its 33*CS words as
// aligned 128-bit shared loads (48-byte lane stride for CS = 3: bank-conflict free).
// A CTA walks kColsRows consecutive V rows; the next row's segment is requested (into registers) before the current
// one is filtered, and two shared-memory buffers alternate (ncu on the one-row version: load -> barrier -> filter with
// nothing to overlap inside a CTA; `long_scoreboard` 15.8 warps per issued instruction).
constexpr int kColsRows = 4;
template <int CS>
__global__ void __launch_bounds__(128)
down4_cols_kernel(const float* __restrict__ V, float* __restrict__ Cout, int RW, int W, int hC, int wC,
                  int row_begin, int row_end, float a, float b) {
  constexpr int kSegPx = 4 * 128 + 2 * kRD + 4;  // pixels staged (a little padding for the last vector)
  constexpr int kSegVec = kSegPx * CS / 4, kPerThread = (kSegVec + 127) / 128;
  __shared__ __align__(16) float seg2[2][kSegPx * CS];
  static_assert((kSegPx * CS) % 4 == 0, "segment must be a whole number of float4");
  const int jy0 = row_begin + blockIdx.y * kColsRows, jy1 = min(jy0 + kColsRows, row_end);
  const long long plane = blockIdx.z;
  const int jx0 = blockIdx.x * 128;
  const int px0 = 4 * (jx0 - kPC) - kRD;
  // interior segment: contiguous and 16-byte aligned (px0 * CS is a multiple of 32 floats) -> 128-bit copies
  const bool fast = px0 >= 0 && px0 + kSegPx <= W && (RW & 3) == 0;
  float4 st[kPerThread];
  auto fetch = [&](int jy) {
    const float4* s4 = reinterpret_cast<const float4*>(V + (plane * hC + jy) * (long long)RW + (long long)px0 * CS);
#pragma unroll
    for (int k = 0; k < kPerThread; ++k) {
      const int i = threadIdx.x + k * 128;
      if (i < kSegVec) st[k] = __ldg(s4 + i);
    }
  };
  if (fast) fetch(jy0);
  const int jx = jx0 + threadIdx.x;
  for (int jy = jy0; jy < jy1; ++jy) {
    float* seg = seg2[(jy - jy0) & 1];
    if (fast) {
#pragma unroll
      for (int k = 0; k < kPerThread; ++k) {
        const int i = threadIdx.x + k * 128;
        if (i < kSegVec) reinterpret_cast<float4*>(seg)[i] = st[k];
      }
    } else {
      const float* src = V + (plane * hC + jy) * (long long)RW;
      for (int i = threadIdx.x; i < kSegPx * CS; i += 128) {
        const int px = i / CS, c = i - px * CS;
        const int xx = min(max(px0 + px, 0), W - 1);
        seg[i] = __ldg(src + xx * CS + c);
      }
    }
    __syncthreads();
    if (fast && jy + 1 < jy1) fetch(jy + 1);
    if (jx < wC) {
      constexpr int kWords = 33 * CS, kVecs = (kWords + 3) / 4;
      const float4* s4 = reinterpret_cast<const float4*>(seg + 4 * threadIdx.x * CS);
      float w[kVecs * 4];
#pragma unroll
      for (int i = 0; i < kVecs; ++i) {
        const float4 t = s4[i];
        w[4 * i] = t.x; w[4 * i + 1] = t.y; w[4 * i + 2] = t.z; w[4 * i + 3] = t.w;
      }
      float acc[CS];
#pragma unroll
      for (int c = 0; c < CS; ++c) acc[c] = 0.f;
#pragma unroll
      for (int t = 0; t <= 2 * kRD; ++t)
#pragma unroll
        for (int c = 0; c < CS; ++c) acc[c] = fmaf(gd(t), w[t * CS + c], acc[c]);
#pragma unroll
      for (int c = 0; c < CS; ++c) Cout[((plane * CS + c) * hC + jy) * (long long)wC + jx] = fmaf(a, acc[c], b);
    }
  }
}

// NHWC (CS = 3), four adjacent outputs per thread.  The windows of adjacent outputs overlap by 29 of 33 pixels, so a
// thread reads its 45 pixels x 3 channels once (34 LDS.128) for 396 FFMAs, where one output per thread needed 25
// LDS.128 per 99 (ncu on that version: shared-memory LSU wavefronts 72 % of peak, `mio_throttle` the top stall).  A
// thread's first pixel is 16 t, i.e. float 48 t: one pad float4 after every 12 makes the lane stride 13 float4
// (conflict-free quarter-warps).  Every output still sums its taps in the order t = 0..32 -- the same bits as
// down4_cols_kernel.  Stores go through a per-warp transpose (a thread's four outputs are consecutive).
constexpr int kCols4Out = 4 * 128;                          // outputs per CTA and row
constexpr int kCols4Px = 4 * kCols4Out + 2 * kRD;           // pixels staged: 4 * 511 + 33 (+ 3: the last float4), in 16-pixel blocks
constexpr int kCols4Vec = kCols4Px * 3 / 4;                 // float4 per segment (a multiple of 12)
constexpr int kCols4Pad = kCols4Vec + kCols4Vec / 12;       // ... with the pads
constexpr int kCols4Stage = (kCols4Vec + 127) / 128;        // float4 per thread and row in flight
static_assert(kCols4Px % 16 == 0 && kCols4Vec % 12 == 0, "segment must be whole 16-pixel (12 float4) blocks");
__global__ void __launch_bounds__(128, 4)  // 128 registers: 0.187 ms on 32 x 4K (142 registers / 3 CTAs: 0.207; 96 / 5: 0.196)
down4_cols4_kernel(const float* __restrict__ V, float* __restrict__ Cout, int RW, int W, int hC, int wC,
                   int row_begin, int row_end, float a, float b) {
  __shared__ __align__(16) float4 seg[kCols4Pad];
  __shared__ float sout[4][32 * 5];
  const int jy0 = row_begin + blockIdx.y * kColsRows, jy1 = min(jy0 + kColsRows, row_end);
  const long long plane = blockIdx.z;
  const int jx0 = blockIdx.x * kCols4Out;
  const int px0 = 4 * (jx0 - kPC) - kRD;
  // float4 i of the segment holds floats px0 * 3 + 4 i ..: a 128-bit load where it lies inside the row (px0 * 3 is a
  // multiple of 4), element-wise edge clamping where it does not (with 512 outputs per CTA EVERY CTA touches a border
  // on 4K frames, so there is no separate "interior" kernel path)
  const bool vec_ok = (RW & 3) == 0;
  float4 st[kCols4Stage];
  auto inside = [&](int i) { const int f0 = px0 * 3 + 4 * i; return vec_ok && f0 >= 0 && f0 + 3 < 3 * W; };
  auto fetch = [&](int jy) {   // the 128-bit loads of the next row, in flight while the current one is filtered
    const float* src = V + (plane * hC + jy) * (long long)RW + px0 * 3;
#pragma unroll
    for (int k = 0; k < kCols4Stage; ++k) {
      const int i = threadIdx.x + k * 128;
      if (i < kCols4Vec && inside(i)) st[k] = __ldg(reinterpret_cast<const float4*>(src) + i);
    }
  };
  fetch(jy0);
  const int t = threadIdx.x, l = t & 31;
  for (int jy = jy0; jy < jy1; ++jy) {
    __syncthreads();  // the previous row's reads of seg are done
#pragma unroll
    for (int k = 0; k < kCols4Stage; ++k) {
      const int i = t + k * 128;
      if (i < kCols4Vec && inside(i)) seg[i + i / 12] = st[k];
    }
    {  // the few float4 that straddle a border of the row: element-wise edge clamp, not prefetched
      const float* src = V + (plane * hC + jy) * (long long)RW;
#pragma unroll 1
      for (int i = t; i < kCols4Vec; i += 128) {
        if (inside(i)) continue;
        float e[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int f = px0 * 3 + 4 * i + j, px = (f + 3 * 64) / 3 - 64, c = f - 3 * px;   // floor division for f >= -192
          e[j] = __ldg(src + min(max(px, 0), W - 1) * 3 + c);
        }
        seg[i + i / 12] = make_float4(e[0], e[1], e[2], e[3]);
      }
    }
    __syncthreads();
    if (jy + 1 < jy1) fetch(jy + 1);
    float acc[4][3];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int c = 0; c < 3; ++c) acc[q][c] = 0.f;
    const float4* s4 = seg + 13 * t;   // float4 12 t of the segment, padded
#pragma unroll
    for (int k = 0; k < 34; ++k) {
      const float4 v4 = s4[k + k / 12];
      const float v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int f = 4 * k + e, u = f / 3, c = f - 3 * u;   // pixel u of the window (0..44), channel c
        if (u <= 44) {
#pragma unroll
          for (int q = 0; q < 4; ++q)
            if (u - 4 * q >= 0 && u - 4 * q <= 2 * kRD) acc[q][c] = fmaf(gd(u - 4 * q), v[e], acc[q][c]);
        }
      }
    }
    // per-warp transpose: lane l holds outputs 4 l .. 4 l + 3 of the warp's 128; store them as four 128-byte lines
    float* wst = sout[t >> 5];
    const int jxw = jx0 + (t & ~31) * 4;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
#pragma unroll
      for (int q = 0; q < 4; ++q) wst[l * 5 + q] = fmaf(a, acc[q][c], b);
      __syncwarp();
      float* dst = Cout + ((plane * 3 + c) * hC + jy) * (long long)wC + jxw;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int o = q * 32 + l;
        if (jxw + o < wC) dst[o] = wst[(o >> 2) * 5 + (o & 3)];
      }
      __syncwarp();
    }
  }
}

// ---- level C -> level B: prefilter + decimate by 4 (same 33-tap Gaussian, in C samples) ---------
// x pass: T[plane][row][jb] = sum_t g[t] C[plane][row][clamp(4(jb-PB)+PC+t-16, 0, wC-1)], rows [r0, r1).
// A CTA stages the edge-clamped C row segment of kDecXTile outputs in shared memory (one pad float after every
// sixteen: thread g reads element 16 g + u at 17 g + ..., conflict-free) and every thread produces four outputs
// from 45 samples instead of 4 x 33; each output still sums its taps in order t = 0..32.
constexpr int kDecXQ = 4;
constexpr int kDecXTile = 128 * kDecXQ;
__global__ void __launch_bounds__(128)
decim4_x_kernel(const float* __restrict__ Cin, float* __restrict__ T, int hC, int wC, int wB, int r0, int r1) {
  constexpr int kSeg = 4 * kDecXTile + 2 * kRD + 16;
  __shared__ float seg[kSeg + kSeg / 16 + 1];
  const long long plane = blockIdx.z;
  const int row = r0 + blockIdx.y;
  const int jb0 = blockIdx.x * kDecXTile;
  const int outs = min(kDecXTile, wB - jb0);
  const float* src = Cin + (plane * hC + row) * (long long)wC;
  const int c0 = 4 * (jb0 - kPB) + kPC - kRD;
  const int need = 4 * outs + 2 * kRD;  // samples the outputs of this CTA reach (rounded up by the loop below)
  for (int i = threadIdx.x; i < need + 12; i += 128) seg[i + (i >> 4)] = __ldg(src + min(max(c0 + i, 0), wC - 1));
  __syncthreads();
  const int g = threadIdx.x;
  if (g * kDecXQ >= outs) return;
  const float* sg = seg + 17 * g;  // element 16 g of the padded segment
  float acc[kDecXQ];
#pragma unroll
  for (int q = 0; q < kDecXQ; ++q) acc[q] = 0.f;
#pragma unroll
  for (int u = 0; u <= 4 * (kDecXQ - 1) + 2 * kRD; ++u) {
    const float v = sg[u + (u >> 4)];
#pragma unroll
    for (int q = 0; q < kDecXQ; ++q)
      if (u - 4 * q >= 0 && u - 4 * q <= 2 * kRD) acc[q] = fmaf(gd(u - 4 * q), v, acc[q]);
  }
  float* dst = T + (plane * hC + row) * (long long)wB + jb0 + g * kDecXQ;
#pragma unroll
  for (int q = 0; q < kDecXQ; ++q)
    if (g * kDecXQ + q < outs) dst[q] = acc[q];
}
// y pass: B[plane][jb][x] = sum_t g[t] T[plane][clamp(4(jb-PB)+PC+t-16, 0, hC-1)][x], B rows [b0, b1).
// One column per lane, kDecQ consecutive B rows per thread: the 4 (kDecQ - 1) + 33 input rows are loaded once
// (coalesced) and every output still sums its taps in order t = 0..32 (bit-identical to one output per thread).
constexpr int kDecQ = 8;
__global__ void __launch_bounds__(128)
decim4_y_kernel(const float* __restrict__ T, float* __restrict__ B, int hC, int hB, int wB, int b0, int b1) {
  const long long plane = blockIdx.z;
  const int xx = blockIdx.x * 128 + threadIdx.x;
  const int jb0 = b0 + blockIdx.y * kDecQ;
  if (xx >= wB || jb0 >= b1) return;
  const float* src = T + plane * (long long)hC * wB + xx;
  const int c0 = 4 * (jb0 - kPB) + kPC - kRD;
  float acc[kDecQ];
#pragma unroll
  for (int q = 0; q < kDecQ; ++q) acc[q] = 0.f;
#pragma unroll
  for (int u = 0; u <= 4 * (kDecQ - 1) + 2 * kRD; ++u) {
    const float v = __ldg(src + (long long)min(max(c0 + u, 0), hC - 1) * wB);
#pragma unroll
    for (int q = 0; q < kDecQ; ++q)
      if (u - 4 * q >= 0 && u - 4 * q <= 2 * kRD) acc[q] = fmaf(gd(u - 4 * q), v, acc[q]);
  }
#pragma unroll
  for (int q = 0; q < kDecQ; ++q)
    if (jb0 + q < b1) B[(plane * hB + jb0 + q) * (long long)wB + xx] = acc[q];
}

// ---- level B -> level C: Gaussian-weighted x4 upsample ------------------------------------------
// C sample jc sits at u = (jc-PC)/4 + PB in B samples; taps j0-4..j0+4 around j0 = round(u) with
// weights exp(-(j-u)^2/2), normalised per phase (phase = (jc-PC) mod 4).
struct UpWeights { float w[4][2 * kTU + 1]; };
__device__ __forceinline__ void up_bc_phase(int jc, int& j0, int& ph) {
  const int d = jc - kPC + 4;  // >= 0
  ph = d & 3;
  j0 = (d >> 2) - 1 + kPB + (ph >= 2 ? 1 : 0);
}
// The four C samples jc = 4q + ph + kPC - 4 (ph = 0..3) of group q share the ten B samples q + kPB - 5 .. q + kPB + 4:
// phases 0, 1 are centred on j0 = q + kPB - 1, phases 2, 3 on j0 + 1.  Both passes therefore load a window of ten
// values once and produce four outputs with compile-time phase (weights become constant-bank FFMA operands).
__device__ __forceinline__ void up_bc_group(const float (&win)[10], const UpWeights& uw, float (&o)[4]) {
#pragma unroll
  for (int ph = 0; ph < 4; ++ph) {
    float acc = 0.f;
#pragma unroll
    for (int t = 0; t <= 2 * kTU; ++t) acc = fmaf(uw.w[ph][t], win[t + (ph >= 2 ? 1 : 0)], acc);
    o[ph] = acc;
  }
}
__host__ __device__ inline int up_bc_x_groups(int wC) { return ((wC - 1 + 4 - kPC) >> 2) + 1; }  // q of the last column + 1
// x pass on B rows [b0, b1): T2[plane][jb][jc] (hB x wC).  One thread per (row, group of 4 output columns).
__global__ void __launch_bounds__(256)
up_bc_x_kernel(const float* __restrict__ SB, float* __restrict__ T2, int hB, int wB, int wC, int b0, int b1,
               const __grid_constant__ UpWeights uw) {
  const long long plane = blockIdx.z;
  const int groups = up_bc_x_groups(wC);
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= (b1 - b0) * groups) return;
  const int jb = b0 + i / groups, q = i % groups;
  const float* src = SB + (plane * hB + jb) * (long long)wB;
  float win[10], o[4];
#pragma unroll
  for (int t = 0; t < 10; ++t) win[t] = __ldg(src + min(max(q + kPB - 5 + t, 0), wB - 1));
  up_bc_group(win, uw, o);
  float* dst = T2 + (plane * hB + jb) * (long long)wC;
#pragma unroll
  for (int ph = 0; ph < 4; ++ph) {
    const int jc = 4 * q + ph + kPC - 4;
    if (jc >= 0 && jc < wC) dst[jc] = o[ph];
  }
}
// y pass for C rows [c0, c1): SC[plane][jc][x] (hC x wC).  One thread per column walks down kUpGroups groups of
// four C rows with the ten-row window in registers: one new T2 sample per four outputs.
constexpr int kUpGroups = 16;
__host__ __device__ inline int up_bc_y_chunks(int c0, int c1) {
  return (((c1 - 1 - kPC + 4) >> 2) - ((c0 - kPC + 4) >> 2)) / kUpGroups + 1;
}
// All stride-16 scales of a call share ONE level-C map M = prod_k S_k^(e_k), e_k = weight_k / weight_ref, because the
// epilogue only needs sum_k weight_k lg S_k = weight_ref lg M.  With equal weights e_k = 1 and M is the plain product.
// The surrounds are smooth at this resolution (sigma >= 40 against a stride of 4), so interpolating M instead of
// every S_k moves the result by < 1e-6 -- and the final kernel upsamples two maps instead of three.  One launch
// upsamples NSRC (1 or 2) scales and writes their product once; further scales multiply in (accumulate = 1).
struct UpSrc { const float* t2[2]; float expo[2]; };
template <int NSRC>
__global__ void __launch_bounds__(128)
up_bc_y_kernel(const __grid_constant__ UpSrc srcs, float* __restrict__ SC, int hB, int hC, int wC, int c0, int c1,
               const __grid_constant__ UpWeights uw, int accumulate) {
  const long long plane = blockIdx.z;
  const int xx = blockIdx.x * 128 + threadIdx.x;
  if (xx >= wC) return;
  const int q_lo = (c0 - kPC + 4) >> 2, q_hi = (c1 - 1 - kPC + 4) >> 2;  // c0 - kPC + 4 >= 0 (kPC = 4)
  const int q0 = q_lo + blockIdx.y * kUpGroups;
  if (q0 > q_hi) return;
  const int q1 = min(q0 + kUpGroups - 1, q_hi);
  const float* src[NSRC];
#pragma unroll
  for (int k = 0; k < NSRC; ++k) src[k] = srcs.t2[k] + plane * (long long)hB * wC + xx;
  float* dst = SC + plane * (long long)hC * wC + xx;
  float win[NSRC][10], o[NSRC][4];
#pragma unroll
  for (int k = 0; k < NSRC; ++k)
#pragma unroll
    for (int t = 1; t < 10; ++t) win[k][t] = __ldg(src[k] + (long long)min(max(q0 + kPB - 6 + t, 0), hB - 1) * wC);
  for (int q = q0; q <= q1; ++q) {
#pragma unroll
    for (int k = 0; k < NSRC; ++k) {
#pragma unroll
      for (int t = 0; t < 9; ++t) win[k][t] = win[k][t + 1];
      win[k][9] = __ldg(src[k] + (long long)min(max(q + kPB + 4, 0), hB - 1) * wC);
      up_bc_group(win[k], uw, o[k]);
    }
#pragma unroll
    for (int ph = 0; ph < 4; ++ph) {
      const int jc = 4 * q + ph + kPC - 4;
      if (jc >= c0 && jc < c1) {
        float v = 1.f;
#pragma unroll
        for (int k = 0; k < NSRC; ++k) v *= srcs.expo[k] != 1.f ? exp2f(srcs.expo[k] * __log2f(o[k][ph])) : o[k][ph];
        float* d = dst + (long long)jc * wC;
        *d = accumulate ? *d * v : v;
      }
    }
  }
}

// ---- K5: x4 Lagrange upsample of every scale's surround + MSRCR epilogue -------------------------
// Thread t of a 128-thread CTA owns coarse column jc0-2+t for the vertical stage and the 4-pixel
// cell t-2 (123 cells = 492 pixels per CTA) for the horizontal stage.  It walks down a strip of
// rows four at a time (one coarse row per step) with the 6 / 4 coarse rows each scale needs held
// in registers, so a step costs one new coarse sample per (scale, channel).  Per step: vertical
// interpolation at the 4 row phases -> shared memory (double-buffered, one barrier per step) ->
// per row: 6 / 4 neighbouring columns from shared memory, horizontal interpolation of the cell's
// 4 pixels, lg2, accumulate, epilogue, 128-bit stores.
struct FinalParams {
  const float* map[3];  // per scale: planar [n][3][hC][wC]; the N6 six-tap scales first
  float w2[3];          // ln2 * weight_k
  float wsum2;          // ln2 * sum of all weights (including earlier passes)
  int hC, wC;
  float beta_ln_alpha, beta_ln2, gain255, offset255, in_scale, in_bias;
  const float* acc_in;  // optional partial sums  sum_k w2_k lg2 S_k  from an earlier pass (image layout)
  float* acc_out;       // if set: write the partial sums instead of the finished pixels
  int no_clip;          // write the un-clipped value (a global-statistics finisher follows)
};

constexpr int kFinalCells = 123;

// log2 of a value >= 1 (I = in_scale * x + in_bias >= in_bias > 0 and its Gaussian averages: the API requires
// in_bias > 0 and the defaults give I in [1, 256]): one MUFU.LG2.  __log2f() wraps the same instruction in a
// denormal-input rescale (FSETP + FMUL + FADD per call: 156 of the 594 instructions of a row of this kernel).
__device__ __forceinline__ float lg2_ge1(float v) {
  float r;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  return r;
}

enum { kIoNhwcVec = 0, kIoNchwVec = 1, kIoScalar = 2 };

// Pull the bytes final_load will read for (row yl, cell X0) into L1 ahead of time: the row loop of the
// final kernel runs at 16 warps/SM (register-bound), too few to hide an HBM round trip per row, and a
// register double buffer would cost 12 more registers (ncu: 34 % of the samples sat on this load).
template <typename T>
__device__ __forceinline__ void final_prefetch(const T* __restrict__ x, const ImgView& v, int io, long long img,
                                               int yl, int X0) {
  if (io == kIoNhwcVec) {
    const T* p = x + img * v.sn + (long long)yl * v.sy + (long long)X0 * 3;
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p + 8));
  } else if (io == kIoNchwVec) {
#pragma unroll
    for (int c = 0; c < 3; ++c)
      asm volatile("prefetch.global.L1 [%0];" ::"l"(x + img * v.sn + c * v.sc + (long long)yl * v.sy + X0));
  }
}

// 4 pixels x 3 channels of one row of the cell: f[px][c].  uint8_t images are read as raw 0..255 values
// (FinalParams.in_scale carries the 1/255) and written as rint(255 * value); uint16_t likewise with 65535.
__device__ __forceinline__ void final_load(const float* __restrict__ x, const ImgView& v, int io, long long img,
                                           int yl, int X0, int W, float (&f)[4][3]) {
  if (io == kIoNhwcVec) {
    const float4* p = reinterpret_cast<const float4*>(x + img * v.sn + (long long)yl * v.sy + (long long)X0 * 3);
    const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    f[0][0] = a.x; f[0][1] = a.y; f[0][2] = a.z; f[1][0] = a.w; f[1][1] = b.x; f[1][2] = b.y;
    f[2][0] = b.z; f[2][1] = b.w; f[2][2] = c.x; f[3][0] = c.y; f[3][1] = c.z; f[3][2] = c.w;
  } else if (io == kIoNchwVec) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(x + img * v.sn + c * v.sc + (long long)yl * v.sy + X0));
      f[0][c] = a.x; f[1][c] = a.y; f[2][c] = a.z; f[3][c] = a.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        f[i][c] = (X0 + i < W) ? __ldg(x + img * v.sn + c * v.sc + (long long)yl * v.sy + (long long)(X0 + i) * v.sx) : 0.f;
  }
}
__device__ __forceinline__ void final_load(const uint8_t* __restrict__ x, const ImgView& v, int io, long long img,
                                           int yl, int X0, int W, float (&f)[4][3]) {
  if (io == kIoNhwcVec) {
    const uchar4* p = reinterpret_cast<const uchar4*>(x + img * v.sn + (long long)yl * v.sy + (long long)X0 * 3);
    const uchar4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    f[0][0] = a.x; f[0][1] = a.y; f[0][2] = a.z; f[1][0] = a.w; f[1][1] = b.x; f[1][2] = b.y;
    f[2][0] = b.z; f[2][1] = b.w; f[2][2] = c.x; f[3][0] = c.y; f[3][1] = c.z; f[3][2] = c.w;
  } else if (io == kIoNchwVec) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const uchar4 a = __ldg(reinterpret_cast<const uchar4*>(x + img * v.sn + c * v.sc + (long long)yl * v.sy + X0));
      f[0][c] = a.x; f[1][c] = a.y; f[2][c] = a.z; f[3][c] = a.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        f[i][c] = (X0 + i < W) ? (float)__ldg(x + img * v.sn + c * v.sc + (long long)yl * v.sy + (long long)(X0 + i) * v.sx) : 0.f;
  }
}
__device__ __forceinline__ void final_load(const uint16_t* __restrict__ x, const ImgView& v, int io, long long img,
                                           int yl, int X0, int W, float (&f)[4][3]) {
  if (io == kIoNhwcVec) {
    const ushort4* p = reinterpret_cast<const ushort4*>(x + img * v.sn + (long long)yl * v.sy + (long long)X0 * 3);
    const ushort4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    f[0][0] = a.x; f[0][1] = a.y; f[0][2] = a.z; f[1][0] = a.w; f[1][1] = b.x; f[1][2] = b.y;
    f[2][0] = b.z; f[2][1] = b.w; f[2][2] = c.x; f[3][0] = c.y; f[3][1] = c.z; f[3][2] = c.w;
  } else if (io == kIoNchwVec) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const ushort4 a = __ldg(reinterpret_cast<const ushort4*>(x + img * v.sn + c * v.sc + (long long)yl * v.sy + X0));
      f[0][c] = a.x; f[1][c] = a.y; f[2][c] = a.z; f[3][c] = a.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        f[i][c] = (X0 + i < W) ? (float)__ldg(x + img * v.sn + c * v.sc + (long long)yl * v.sy + (long long)(X0 + i) * v.sx) : 0.f;
  }
}
__device__ __forceinline__ void final_store(float* __restrict__ o, const ImgView& v, int io, long long img, int yl,
                                            int X0, int W, const float (&f)[4][3]) {
  if (io == kIoNhwcVec) {
    float4* p = reinterpret_cast<float4*>(o + img * v.sn + (long long)yl * v.sy + (long long)X0 * 3);
    p[0] = make_float4(f[0][0], f[0][1], f[0][2], f[1][0]);
    p[1] = make_float4(f[1][1], f[1][2], f[2][0], f[2][1]);
    p[2] = make_float4(f[2][2], f[3][0], f[3][1], f[3][2]);
  } else if (io == kIoNchwVec) {
#pragma unroll
    for (int c = 0; c < 3; ++c)
      *reinterpret_cast<float4*>(o + img * v.sn + c * v.sc + (long long)yl * v.sy + X0) =
          make_float4(f[0][c], f[1][c], f[2][c], f[3][c]);
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        if (X0 + i < W) o[img * v.sn + c * v.sc + (long long)yl * v.sy + (long long)(X0 + i) * v.sx] = f[i][c];
  }
}
__device__ __forceinline__ unsigned char to_u8(float v01) { return (unsigned char)__float2uint_rn(255.f * v01); }
__device__ __forceinline__ void final_store(uint8_t* __restrict__ o, const ImgView& v, int io, long long img, int yl,
                                            int X0, int W, const float (&f)[4][3]) {
  if (io == kIoNhwcVec) {
    uchar4* p = reinterpret_cast<uchar4*>(o + img * v.sn + (long long)yl * v.sy + (long long)X0 * 3);
    p[0] = make_uchar4(to_u8(f[0][0]), to_u8(f[0][1]), to_u8(f[0][2]), to_u8(f[1][0]));
    p[1] = make_uchar4(to_u8(f[1][1]), to_u8(f[1][2]), to_u8(f[2][0]), to_u8(f[2][1]));
    p[2] = make_uchar4(to_u8(f[2][2]), to_u8(f[3][0]), to_u8(f[3][1]), to_u8(f[3][2]));
  } else if (io == kIoNchwVec) {
#pragma unroll
    for (int c = 0; c < 3; ++c)
      *reinterpret_cast<uchar4*>(o + img * v.sn + c * v.sc + (long long)yl * v.sy + X0) =
          make_uchar4(to_u8(f[0][c]), to_u8(f[1][c]), to_u8(f[2][c]), to_u8(f[3][c]));
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        if (X0 + i < W) o[img * v.sn + c * v.sc + (long long)yl * v.sy + (long long)(X0 + i) * v.sx] = to_u8(f[i][c]);
  }
}

__device__ __forceinline__ unsigned short to_u16(float v01) { return (unsigned short)__float2uint_rn(65535.f * v01); }
__device__ __forceinline__ void final_store(uint16_t* __restrict__ o, const ImgView& v, int io, long long img, int yl,
                                            int X0, int W, const float (&f)[4][3]) {
  if (io == kIoNhwcVec) {
    ushort4* p = reinterpret_cast<ushort4*>(o + img * v.sn + (long long)yl * v.sy + (long long)X0 * 3);
    p[0] = make_ushort4(to_u16(f[0][0]), to_u16(f[0][1]), to_u16(f[0][2]), to_u16(f[1][0]));
    p[1] = make_ushort4(to_u16(f[1][1]), to_u16(f[1][2]), to_u16(f[2][0]), to_u16(f[2][1]));
    p[2] = make_ushort4(to_u16(f[2][2]), to_u16(f[3][0]), to_u16(f[3][1]), to_u16(f[3][2]));
  } else if (io == kIoNchwVec) {
#pragma unroll
    for (int c = 0; c < 3; ++c)
      *reinterpret_cast<ushort4*>(o + img * v.sn + c * v.sc + (long long)yl * v.sy + X0) =
          make_ushort4(to_u16(f[0][c]), to_u16(f[1][c]), to_u16(f[2][c]), to_u16(f[3][c]));
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        if (X0 + i < W) o[img * v.sn + c * v.sc + (long long)yl * v.sy + (long long)(X0 + i) * v.sx] = to_u16(f[i][c]);
  }
}

// x / out point at image row in_row0 (row-tiled shards); rows [row_begin, row_end) are produced,
// row_begin a multiple of 4, rows_per_strip a multiple of 4.
// kEqW: all scale weights are equal (the default 1/3, 1/3, 1/3), so sum_k w lg2 S_k = w lg2(prod_k S_k): one MUFU.LG2
// per pixel-channel instead of one per scale (13 -> 7 per pixel; the XU pipe, 16 lanes/clk/SM, was ~50 % of the
// kernel's time).  S_k <= in_scale + in_bias, so the product of up to three stays far inside fp32 range.
// kIo >= 0: the I/O layout (kIoNhwcVec / kIoNchwVec) is a compile-time constant and there are no partial sums from /
// for another pass (the default call: one launch, vector I/O) -- the other layouts' load / store code and the
// acc_in / acc_out branches leave the row loop.  kIo = -1: everything decided at run time.
template <int N6, int N4, typename TIO, bool kEqW, int kIo = -1>
__global__ void __launch_bounds__(128, 5)
msrcr_final_kernel(const TIO* __restrict__ x, TIO* __restrict__ out, ImgView v, int io_rt, int W, int in_row0,
                   int row_begin, int row_end, int rows_per_strip, const __grid_constant__ FinalParams p) {
  constexpr int NV = 3 * (N6 + N4);
  const int io = kIo >= 0 ? kIo : io_rt;
  constexpr bool kAcc = kIo < 0;  // partial sums in / out possible
  __shared__ float s_v[2][4][NV][128];
  const int t = threadIdx.x;
  const long long img = blockIdx.z;
  const int cell0 = blockIdx.x * kFinalCells;
  const int jcv = min(max(kPC + cell0 - 2 + t, 0), p.wC - 1);  // coarse column of the vertical duty
  const int cell = t - 2;
  const int X0 = 4 * (cell0 + cell);
  const bool h_on = cell >= 0 && cell < kFinalCells && X0 < W;
  const int ys = row_begin + blockIdx.y * rows_per_strip;
  const int ye = min(ys + rows_per_strip, row_end);
  if (ys >= ye) return;
  const int hC = p.hC, wC = p.wC;

  const float* m6[N6 > 0 ? N6 : 1];
  const float* m4[N4 > 0 ? N4 : 1];
#pragma unroll
  for (int k = 0; k < N6; ++k) m6[k] = p.map[k] + img * 3ll * hC * wC + jcv;
#pragma unroll
  for (int k = 0; k < N4; ++k) m4[k] = p.map[N6 + k] + img * 3ll * hC * wC + jcv;
  auto ld = [&](const float* m, int c, int row) { return __ldg(m + ((long long)c * hC + min(row, hC - 1)) * wC); };

  // windows: win6 rows j0-2..j0+3, win4 rows j0-1..j0+2 of the current step
  float win6[N6 > 0 ? N6 : 1][3][6], win4[N4 > 0 ? N4 : 1][3][4];
  const int jfirst = ys / 4 + kPC;
#pragma unroll
  for (int k = 0; k < N6; ++k)
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int i = 0; i < 6; ++i) win6[k][c][i] = ld(m6[k], c, jfirst - 2 + i);
#pragma unroll
  for (int k = 0; k < N4; ++k)
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int i = 0; i < 4; ++i) win4[k][c][i] = ld(m4[k], c, jfirst - 1 + i);

  const int nsteps = (ye - ys + 3) / 4;
  if (h_on && ys + 1 < ye) final_prefetch(x, v, io, img, ys + 1 - in_row0, X0);
  for (int s = 0; s < nsteps; ++s) {
    const int Y0 = ys + 4 * s, j0 = Y0 / 4 + kPC;
    float (*sv)[NV][128] = s_v[s & 1];
    // prefetch the next step's new coarse rows
    float n6[N6 > 0 ? N6 : 1][3], n4[N4 > 0 ? N4 : 1][3];
#pragma unroll
    for (int k = 0; k < N6; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) n6[k][c] = ld(m6[k], c, j0 + 4);
#pragma unroll
    for (int k = 0; k < N4; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) n4[k][c] = ld(m4[k], c, j0 + 3);
    // vertical stage
#pragma unroll
    for (int k = 0; k < N6; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float* w = win6[k][c];
        sv[0][k * 3 + c][t] = w[2];
#pragma unroll
        for (int ph = 1; ph < 4; ++ph) {
          float a = kL6[ph - 1][0] * w[0];
#pragma unroll
          for (int i = 1; i < 6; ++i) a = fmaf(kL6[ph - 1][i], w[i], a);
          sv[ph][k * 3 + c][t] = a;
        }
      }
#pragma unroll
    for (int k = 0; k < N4; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float* w = win4[k][c];
        sv[0][(N6 + k) * 3 + c][t] = w[1];
#pragma unroll
        for (int ph = 1; ph < 4; ++ph) {
          float a = kL4[ph - 1][0] * w[0];
#pragma unroll
          for (int i = 1; i < 4; ++i) a = fmaf(kL4[ph - 1][i], w[i], a);
          sv[ph][(N6 + k) * 3 + c][t] = a;
        }
      }
    __syncthreads();
    // horizontal stage + epilogue, one image row per phase
    if (h_on) {
#pragma unroll 1
      for (int ph = 0; ph < 4; ++ph) {
        const int Y = Y0 + ph;
        if (Y >= ye) break;
        const int yl = Y - in_row0;
        float f[4][3], acc[4][3];
        if (Y + 2 < ye) final_prefetch(x, v, io, img, yl + 2, X0);  // two rows ahead: one row is shorter than the DRAM latency
        final_load(x, v, io, img, yl, X0, W, f);
        if (kAcc && p.acc_in) {
          final_load(p.acc_in, v, io, img, yl, X0, W, acc);
        } else {
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int c = 0; c < 3; ++c) acc[i][c] = 0.f;
        }
        float prod[4][3];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int c = 0; c < 3; ++c) prod[i][c] = 1.f;
        auto add_scale = [&](int q, int c, int k, float sval) {
          if constexpr (kEqW) prod[q][c] *= sval;
          else acc[q][c] = fmaf(p.w2[k], lg2_ge1(sval), acc[q][c]);
        };
#pragma unroll
        for (int k = 0; k < N6; ++k)
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            float w[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) w[i] = sv[ph][k * 3 + c][cell + i];
            add_scale(0, c, k, w[2]);
#pragma unroll
            for (int q = 1; q < 4; ++q) {
              float a = kL6[q - 1][0] * w[0];
#pragma unroll
              for (int i = 1; i < 6; ++i) a = fmaf(kL6[q - 1][i], w[i], a);
              add_scale(q, c, k, a);
            }
          }
#pragma unroll
        for (int k = 0; k < N4; ++k)
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            float w[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) w[i] = sv[ph][(N6 + k) * 3 + c][cell + 1 + i];
            add_scale(0, c, N6 + k, w[1]);
#pragma unroll
            for (int q = 1; q < 4; ++q) {
              float a = kL4[q - 1][0] * w[0];
#pragma unroll
              for (int i = 1; i < 4; ++i) a = fmaf(kL4[q - 1][i], w[i], a);
              add_scale(q, c, N6 + k, a);
            }
          }
        if constexpr (kEqW) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int c = 0; c < 3; ++c) acc[i][c] = fmaf(p.w2[0], lg2_ge1(prod[i][c]), acc[i][c]);
        }
        if (kAcc && p.acc_out) {
          final_store(p.acc_out, v, io, img, yl, X0, W, acc);
        } else {
          float o[4][3];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float I[3], l[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              I[c] = fmaf(p.in_scale, f[i][c], p.in_bias);
              l[c] = lg2_ge1(I[c]);
            }
            const float lsum = lg2_ge1(I[0] + I[1] + I[2]);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              const float R = fmaf(p.wsum2, l[c], -acc[i][c]);
              const float Cc = fmaf(p.beta_ln2, l[c] - lsum, p.beta_ln_alpha);
              const float val = fmaf(p.gain255 * Cc, R, p.offset255);
              o[i][c] = p.no_clip ? val : __saturatef(val);
            }
          }
          final_store(out, v, io, img, yl, X0, W, o);
        }
      }
    }
    // slide the windows down one coarse row
#pragma unroll
    for (int k = 0; k < N6; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
#pragma unroll
        for (int i = 0; i < 5; ++i) win6[k][c][i] = win6[k][c][i + 1];
        win6[k][c][5] = n6[k][c];
      }
#pragma unroll
    for (int k = 0; k < N4; ++k)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
#pragma unroll
        for (int i = 0; i < 3; ++i) win4[k][c][i] = win4[k][c][i + 1];
        win4[k][c][3] = n4[k][c];
      }
  }
}

}  // namespace
}  // namespace llie
