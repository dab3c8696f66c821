// Generic edge-clamped 1-D FIR kernels (fp32) shared by the exact MSRCR path (msrcr_direct.cu)
// and the low-resolution stages of the pyramid path (msrcr_pyramid.cu).
#pragma once
#include "llie_common.cuh"

namespace llie {
namespace {

constexpr int kFold = 64;  // taps accumulated locally before folding into the running sum

// FIR along x.  Data is rows of W pixels with C interleaved channels (C=3: NHWC image rows;
// C=1: NCHW plane rows).  Each thread produces 4 consecutive pixels of one channel with a
// sliding set of 4 weights, reading the clamped input segment from shared memory.
// Row addressing: blockIdx.x = p * rows_per_plane + i handles buffer row p * plane_rows + row_lo + i
// (a band of rows_per_plane rows in each plane of plane_rows rows).  All rows of a buffer:
// rows_per_plane = gridDim.x, plane_rows = 0, row_lo = 0.
template <bool kAffine>
__global__ void __launch_bounds__(128) fir_x_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                    const float* __restrict__ taps, int r, int W, int C,
                                                    float a, float b, int rows_per_plane, int plane_rows, int row_lo) {
  extern __shared__ float seg[];
  const int TX = 128;
  const int x0 = blockIdx.y * TX;
  const int pl = blockIdx.x / rows_per_plane;
  const long long row = (long long)pl * plane_rows + row_lo + (blockIdx.x - pl * rows_per_plane);
  const float* src = in + row * (long long)W * C;
  const int seg_px = TX + 2 * r + 3;
  for (int i = threadIdx.x; i < seg_px * C; i += blockDim.x) {
    int px = i / C, c = i - px * C;
    int x = min(max(x0 - r + px, 0), W - 1);
    float v = src[(long long)x * C + c];
    seg[i] = kAffine ? fmaf(a, v, b) : v;
  }
  __syncthreads();
  const int groups = TX / 4;
  for (int t = threadIdx.x; t < groups * C; t += blockDim.x) {
    int g = t / C, c = t - g * C;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    float k1 = 0.f, k2 = 0.f, k3 = 0.f;
    const float* s = seg + (g * 4) * C + c;
    const int n_steps = 2 * r + 4;
    for (int u0 = 0; u0 < n_steps; u0 += kFold) {
      float p[4] = {0.f, 0.f, 0.f, 0.f};
      int u1 = min(u0 + kFold, n_steps);
      for (int u = u0; u < u1; ++u) {
        float v = s[u * C];
        float k0 = (u <= 2 * r) ? taps[u] : 0.f;
        p[0] = fmaf(k0, v, p[0]);
        p[1] = fmaf(k1, v, p[1]);
        p[2] = fmaf(k2, v, p[2]);
        p[3] = fmaf(k3, v, p[3]);
        k3 = k2; k2 = k1; k1 = k0;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[j] += p[j];
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int x = x0 + g * 4 + j;
      if (x < W) out[(row * W + x) * C + c] = acc[j];
    }
  }
}

// FIR along y on `planes` planes stored as in_rows x RW (input) and out_plane_rows x RW (output).
// Rows are addressed by their index in a level of H rows: buffer row b of the input holds level
// row in_row0 + b, and outputs are produced for level rows out_row0 .. out_row0+out_rows-1, stored
// at output buffer rows out_buf_row0 .. ; taps that fall outside [0, H) are clamped (the caller
// guarantees the clamped rows are in the buffer).  One image: in_row0 = out_row0 = out_buf_row0 = 0,
// in_rows = out_rows = out_plane_rows = H.  4 consecutive rows per thread.
__global__ void __launch_bounds__(128) fir_y_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                    const float* __restrict__ taps, int r, int H, int RW,
                                                    int in_row0, int in_rows, int out_row0, int out_rows,
                                                    int out_buf_row0, int out_plane_rows) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= RW) return;
  const int yb = blockIdx.y * 4;           // first output buffer row of this thread
  const int y0 = out_row0 + yb;            // its level row
  const long long plane = blockIdx.z;
  const float* src = in + plane * (long long)in_rows * RW + j;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  float k1 = 0.f, k2 = 0.f, k3 = 0.f;
  const int n_steps = 2 * r + 4;
  for (int u0 = 0; u0 < n_steps; u0 += kFold) {
    float p[4] = {0.f, 0.f, 0.f, 0.f};
    int u1 = min(u0 + kFold, n_steps);
    for (int u = u0; u < u1; ++u) {
      int y = min(max(y0 - r + u, 0), H - 1) - in_row0;
      float v = __ldg(src + (long long)y * RW);
      float k0 = (u <= 2 * r) ? taps[u] : 0.f;
      p[0] = fmaf(k0, v, p[0]);
      p[1] = fmaf(k1, v, p[1]);
      p[2] = fmaf(k2, v, p[2]);
      p[3] = fmaf(k3, v, p[3]);
      k3 = k2; k2 = k1; k1 = k0;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[q] += p[q];
  }
  float* dst = out + (plane * (long long)out_plane_rows + out_buf_row0) * RW + j;
#pragma unroll
  for (int q = 0; q < 4; ++q)
    if (yb + q < out_rows) dst[(long long)(yb + q) * RW] = acc[q];
}

// ---- planar (C = 1) variants for the low-resolution levels of the cascade ------------------------
// Eight outputs per thread and the tap loop unrolled by eight, so that the sliding taps sit in eight registers
// whose roles rotate at compile time: per eight input samples 8 loads + 8 tap loads + 64 FFMA (the generic
// kernels above spend 3 in 4 instructions on bookkeeping, and fir_x_kernel keeps only 32 of 128 threads busy at
// C = 1).  Every output accumulates its taps strictly in order t = 0 .. 2r, whatever its position in the
// thread's group of eight, so results do not depend on how rows / columns are grouped (row-tiled == whole).
constexpr int kFir8 = 8;
// Taps as a by-value kernel argument (constant bank): the tap of a step is then a uniform LDC instead of a
// shared-memory load, which halves the load-store-unit work of the inner loop (8 LDS/LDG + 8 tap LDS per 64 FFMA
// kept the LSU as busy as the FMA pipes).  Filters longer than kFirParamTaps fall back to taps in shared memory.
constexpr int kFirParamTaps = 160;
struct FirTaps { float t[kFirParamTaps]; };
constexpr int kFirX8Tile = 128 * kFir8;  // pixels per CTA of fir_x8_kernel
__host__ __device__ inline int fir8_steps(int r) { return (2 * r + kFir8 + 7) & ~7; }
// Row addressing as fir_x_kernel, for the flat row index blockIdx.x * kFirXRows + i < total_rows.  Shared memory: two
// buffers of the clamped input segment, one pad float after every eight (thread g reads element 8 g + u: bank
// 9 g + ..., conflict-free), then the zero-padded taps.  A CTA walks kFirXRows rows: the next row's samples are
// loaded into registers before the current row is filtered (ncu on the one-row version: load -> barrier -> compute
// with nothing to overlap inside a CTA; issue slots 60 % busy at 90 % occupancy).
constexpr int kFirXRows = 4;
constexpr int kFirXStage = (kFirX8Tile + 160 + 127) / 128;  // staged floats per thread and row (tile + longest filter)
__host__ __device__ inline size_t fir_x8_smem(int r) {
  const int seg = kFirX8Tile + fir8_steps(r);
  return (size_t)(2 * (seg + (seg >> 3) + 1) + 4 * 32 * 9 + fir8_steps(r)) * sizeof(float);
}
template <bool kParamTaps, int kSteps>
__global__ void __launch_bounds__(128) fir_x8_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                     const float* __restrict__ taps, int r, int W,
                                                     int rows_per_plane, int plane_rows, int row_lo, int total_rows,
                                                     const __grid_constant__ FirTaps ft) {
  extern __shared__ float seg_all[];
  const int x0 = blockIdx.y * kFirX8Tile;
  const int n_steps = kSteps > 0 ? kSteps : fir8_steps(r);
  const int tile = min(kFirX8Tile, (W - x0 + 7) & ~7);
  const int seg_px = tile + n_steps;
  const int seg_stride = kFirX8Tile + n_steps + ((kFirX8Tile + n_steps) >> 3) + 1;
  float* sout = seg_all + 2 * seg_stride;  // 4 warps x 32 threads x (8 outputs + 1 pad)
  float* stap = sout + 4 * 32 * 9;
  if constexpr (!kParamTaps)
    for (int i = threadIdx.x; i < n_steps; i += blockDim.x) stap[i] = (i <= 2 * r) ? __ldg(taps + i) : 0.f;
  const int ri0 = blockIdx.x * kFirXRows, ri1 = min(ri0 + kFirXRows, total_rows);
  auto row_of = [&](int ri) -> long long {
    const int pl = ri / rows_per_plane;
    return (long long)pl * plane_rows + row_lo + (ri - pl * rows_per_plane);
  };
  float st[kFirXStage];
  // Samples i < 128 kFirXStage of row ri go through registers (stored to shared memory one barrier later).  Only
  // filters too long for the by-value taps can have a longer segment (tile + n_steps > 1280): that tail is written
  // straight into row ri's buffer, which is free -- its previous user, row ri - 2, was finished by every thread
  // before the barrier of row ri - 1, and fetch(ri) runs after that barrier.
  auto fetch = [&](int ri) {
    const float* src = in + row_of(ri) * (long long)W;
#pragma unroll
    for (int k = 0; k < kFirXStage; ++k) {
      const int i = threadIdx.x + k * 128;
      if (i < seg_px) st[k] = __ldg(src + min(max(x0 - r + i, 0), W - 1));
    }
    if constexpr (!kParamTaps) {
      float* tail = seg_all + ((ri - ri0) & 1) * seg_stride;
      for (int i = threadIdx.x + kFirXStage * 128; i < seg_px; i += 128)
        tail[i + (i >> 3)] = __ldg(src + min(max(x0 - r + i, 0), W - 1));
    }
  };
  fetch(ri0);
  const int g = threadIdx.x;
  for (int ri = ri0; ri < ri1; ++ri) {
    float* seg = seg_all + ((ri - ri0) & 1) * seg_stride;
#pragma unroll
    for (int k = 0; k < kFirXStage; ++k) {
      const int i = threadIdx.x + k * 128;
      if (i < seg_px) seg[i + (i >> 3)] = st[k];
    }
    __syncthreads();
    if (ri + 1 < ri1) fetch(ri + 1);
    if ((g & ~31) * kFir8 < tile) {  // whole warps: the store below is warp-cooperative
      float acc[kFir8], k[kFir8];
#pragma unroll
      for (int q = 0; q < kFir8; ++q) acc[q] = 0.f, k[q] = 0.f;
      const float* sg = seg + g * 9;  // element 8 g of the padded segment
      auto block = [&](int u0) {
        const float* sb = sg + u0 + (u0 >> 3);
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          const float v = sb[s];
          k[s] = kParamTaps ? ft.t[u0 + s] : stap[u0 + s];
#pragma unroll
          for (int q = 0; q < kFir8; ++q) acc[q] = fmaf(k[(s - q) & 7], v, acc[q]);
        }
      };
      if constexpr (kSteps > 0) {  // compile-time tap count: taps become constant-bank FFMA operands, offsets immediates
#pragma unroll
        for (int u0 = 0; u0 < kSteps; u0 += 8) block(u0);
      } else {
        for (int u0 = 0; u0 < n_steps; u0 += 8) block(u0);
      }
      // a thread's eight outputs are consecutive, so storing them directly would touch 32 sectors per warp store
      // (4 bytes each): transpose through the warp's strip of shared memory and store whole 128-byte lines instead
      float* wst = sout + (threadIdx.x >> 5) * (32 * 9);
      const int l = threadIdx.x & 31;
#pragma unroll
      for (int q = 0; q < kFir8; ++q) wst[l * 9 + q] = acc[q];
      __syncwarp();
      float* dst = out + row_of(ri) * (long long)W + x0 + (threadIdx.x & ~31) * kFir8;
#pragma unroll
      for (int q = 0; q < kFir8; ++q) {
        const int o = q * 32 + l;  // output o of this warp's 256: thread o / 8, slot o % 8
        if (x0 + (threadIdx.x & ~31) * kFir8 + o < W) dst[o] = wst[(o >> 3) * 9 + (o & 7)];
      }
      __syncwarp();
    }
  }
}

// Row addressing as fir_y_kernel; 8 consecutive rows per thread, one column per lane (coalesced).
// kSteps > 0: the number of 8-tap steps is a compile-time constant, so the tap loop unrolls completely and every tap
// is a constant-bank operand of its FFMA (ncu on the dynamic loop: 1000 instructions per thread for 320 FFMAs --
// the tap fetch and the 64-bit clamped row address of every load cost more issue slots than the arithmetic; issue
// slots 81 % busy).  Threads whose whole window lies inside the level (all but the border rows) step a pointer.
template <bool kParamTaps, int kSteps>
__global__ void __launch_bounds__(128) fir_y8_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                     const float* __restrict__ taps, int r, int H, int RW,
                                                     int in_row0, int in_rows, int out_row0, int out_rows,
                                                     int out_buf_row0, int out_plane_rows,
                                                     const __grid_constant__ FirTaps ft) {
  extern __shared__ float stap[];
  const int n_steps = kSteps > 0 ? kSteps : fir8_steps(r);
  if constexpr (!kParamTaps) {
    for (int i = threadIdx.x; i < n_steps; i += blockDim.x) stap[i] = (i <= 2 * r) ? __ldg(taps + i) : 0.f;
    __syncthreads();
  }
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= RW) return;
  const int yb = blockIdx.y * kFir8;       // first output buffer row of this thread
  const int y0 = out_row0 + yb;            // its level row
  const long long plane = blockIdx.z;
  const float* src = in + plane * (long long)in_rows * RW + j;
  float acc[kFir8], k[kFir8], v[8], vn[8];
#pragma unroll
  for (int q = 0; q < kFir8; ++q) acc[q] = 0.f, k[q] = 0.f;
  const bool interior = y0 - r >= 0 && y0 - r + n_steps - 1 <= H - 1;
  const float* p = src + (long long)(y0 - r - in_row0) * RW;
  // the eight samples of the NEXT block are requested before the 64 FFMAs of the current one
  auto load8 = [&](float (&d)[8], int u0) {
    if (interior) {
#pragma unroll
      for (int s = 0; s < 8; ++s) { d[s] = __ldg(p); p += RW; }
    } else {
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        const int y = min(max(y0 - r + u0 + s, 0), H - 1) - in_row0;
        d[s] = __ldg(src + (long long)y * RW);
      }
    }
  };
  auto block = [&](int u0) {
    if (u0 + 8 < n_steps) load8(vn, u0 + 8);
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      k[s] = kParamTaps ? ft.t[u0 + s] : stap[u0 + s];
#pragma unroll
      for (int q = 0; q < kFir8; ++q) acc[q] = fmaf(k[(s - q) & 7], v[s], acc[q]);
    }
#pragma unroll
    for (int s = 0; s < 8; ++s) v[s] = vn[s];
  };
  load8(v, 0);
  if constexpr (kSteps > 0) {
#pragma unroll
    for (int u0 = 0; u0 < kSteps; u0 += 8) block(u0);
  } else {
    for (int u0 = 0; u0 < n_steps; u0 += 8) block(u0);
  }
  float* dst = out + (plane * (long long)out_plane_rows + out_buf_row0) * RW + j;
#pragma unroll
  for (int q = 0; q < kFir8; ++q)
    if (yb + q < out_rows) dst[(long long)(yb + q) * RW] = acc[q];
}

}  // namespace
}  // namespace llie
