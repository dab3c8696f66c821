// Zero-DCE, bf16 tensor-core path (config 2): DCE-Net conv2..conv7 as implicit-GEMM 3x3
// convolutions on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM), conv1 as a
// direct CUDA-core convolution, and tanh + the 8 curve iterations fused into conv7's epilogue
// (north_star kernels (2) and (3)).  No reference file exists to cite
// (/root/reference/README.md:1-2).
//
// Formulation.  Pixels are the MMA's M dimension: one MMA covers 128 consecutive pixels of one
// image row.  Activations live in HBM as [n][y][chunk=C/8][x+1][8 ch] bf16 with zero pad columns,
// so one input row of a 128-pixel strip is C/8 contiguous runs of 130 pixels x 16 B; a 1-D bulk
// async copy (TMA engine) drops each run into shared memory exactly in the K-major no-swizzle
// layout tcgen05 wants (row = pixel, 16 B apart; K halves one run apart).  A horizontal tap dx
// is then just the A descriptor's start address + 16*dx bytes -- no im2col, no data movement.
// The three vertical taps are stacked in N: B = [dy=+1 | dy=0 | dy=-1] x 32 output channels
// (N = 96), and the TMEM accumulators of consecutive output rows sit in consecutive 32-column
// slots, so ONE N=96 MMA on input row y' accumulates into output rows y'-1, y', y'+1 at once.
// N = 96 instead of 32 cuts the A-operand shared-memory traffic per MAC by 3.
// Bias is free: the epilogue re-initialises a drained slot with the bias vector (tcgen05.st).
//
// Roles per CTA (persistent, 1 CTA/SM, 320 threads): warp 0 = copy producer, warp 1 = MMA issuer
// (+ TMEM allocator), warps 2..9 = epilogue (two groups of 4 warps alternate output rows; each
// warp owns the TMEM lane quadrant warp%4).  A work unit is (image, 128-px strip, row segment);
// input rows stream through an 8-stage ring, output rows through a 16-slot TMEM ring.
#include "dce_weights.cuh"
#include "sm100_ptx.cuh"

#include <cstring>

namespace llie {

namespace {
using namespace ptx;

constexpr int kTileX = 128;                 // pixels per MMA (M)
constexpr int kRowPx = kTileX + 2;          // staged pixels per input row (1-px halo each side)
constexpr int kChunkBytes = kRowPx * 16;    // 2080: one 8-channel chunk of one staged row
constexpr int kFirstRowBytes = (kRowPx + 2) * 16;  // conv1 stages 132 px: its dx=2 MMA's zero-weight K half reads px+3
constexpr int kSlots = 16;                  // TMEM ring: 16 output rows x 32 fp32 columns = 512
constexpr int kStages = 8;                  // input-row ring
constexpr int kBTile = 96 * 16 * 2;         // 3072 B: weight tile of one (dx, k16): N=96 x K=16
constexpr int kBPlane = 96 * 16;            // 1536 B between the two K halves of a weight tile
constexpr int kEpiWarps = 8;
constexpr int kThreads = (2 + kEpiWarps) * 32;

struct ConvArgs {
  const __nv_bfloat16* in_a;  // [n][H][4][Wp][8]
  const __nv_bfloat16* in_b;  // second source of a concat layer (channels 32..63) or nullptr
  const __nv_bfloat16* wpk;   // 3 * Cin/16 weight tiles
  const float* bias;          // 32 floats
  __nv_bfloat16* out;         // [n][H][4][Wp][8]; unused by the last layer
  const float* x;             // last layer: the caller's image / output / optional maps
  float* y;
  float* maps;
  ImgView vx, vm;
  int iters;
  float sign;
  int N, H, W, Wp, strips, segs, seg_rows, units;
};

__device__ __forceinline__ uint32_t pack_bf16_relu(uint32_t a, uint32_t b) {
  __nv_bfloat162 p = __floats2bfloat162_rn(fmaxf(__uint_as_float(a), 0.f), fmaxf(__uint_as_float(b), 0.f));
  return *reinterpret_cast<uint32_t*>(&p);
}
__device__ __forceinline__ float fast_tanh(float v) {
  // 1 - 2/(e^{2v}+1); |error| ~1e-6, saturates cleanly for large |v|
  float e = __expf(2.f * v);
  return 1.f - __fdividef(2.f, e + 1.f);
}

// FIRST (conv1, 3 -> 32): the input is ONE 8-channel chunk (RGB + 5 zero channels) and the two K
// halves of an MMA are two ADJACENT PIXELS of it (descriptor LBO = 16 B), so K = 16 covers the taps
// dx = 0,1 and a second MMA covers dx = 2 (+ a zero-weight half): 2 MMAs per input row.
template <int CHUNKS, bool LAST, bool FIRST = false>
__global__ void __launch_bounds__(kThreads, 1) dce_conv_tc_kernel(const __grid_constant__ ConvArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  static_assert(!FIRST || (CHUNKS == 1 && !LAST), "conv1 variant is single-chunk");
  constexpr int kWBytes = FIRST ? 2 * kBTile : 3 * (CHUNKS / 2) * kBTile;
  constexpr int kRowBytes = FIRST ? kFirstRowBytes : kChunkBytes;  // bytes of one staged chunk row
  constexpr int kStageBytes = CHUNKS * kRowBytes;
  uint8_t* s_w = smem;
  uint8_t* s_a = smem + kWBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(s_a + kStages * kStageBytes);
  uint64_t* empty = full + kStages;
  uint64_t* acc_ready = empty + kStages;
  uint64_t* acc_free = acc_ready + kSlots;
  uint64_t* wbar = acc_free + kSlots;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(wbar + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int s = 0; s < kSlots; ++s) { mbar_init(&acc_ready[s], 1); mbar_init(&acc_free[s], 4); }
    mbar_init(wbar, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(s_tmem, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *s_tmem;

  const int H = a.H;
  auto decode = [&](int u, int& n, int& sx, int& ys, int& ye) {
    int seg = u % a.segs;
    int t = u / a.segs;
    sx = t % a.strips;
    n = t / a.strips;
    ys = seg * a.seg_rows;
    ye = min(ys + a.seg_rows, H);
  };

  if (warp == 0) {
    // ---------------- producer: one bulk copy per (input row, 8-channel chunk) ----------------
    // The whole warp walks the loop (warp-uniform control flow keeps addresses in uniform
    // registers); one elected lane issues the copies.
    if (elect_one()) {
      mbar_expect_tx(wbar, kWBytes);
      bulk_g2s(s_w, a.wpk, kWBytes, wbar);
    }
    __syncwarp();
    uint32_t irow = 0;
    for (int u = blockIdx.x; u < a.units; u += gridDim.x) {
      int n, sx, ys, ye;
      decode(u, n, sx, ys, ye);
      const int yi0 = max(ys - 1, 0), yi1 = min(ye, H - 1);
      const size_t row_stride = (size_t)(FIRST ? 1 : 4) * a.Wp * 8;  // elements between consecutive rows of one source
      size_t off = ((size_t)n * H + yi0) * row_stride + (size_t)sx * kTileX * 8;
      for (int yi = yi0; yi <= yi1; ++yi, off += row_stride, ++irow) {
        const uint32_t stage = irow & (kStages - 1);
        mbar_wait(&empty[stage], ((irow / kStages) & 1) ^ 1);
        if (elect_one()) {
          mbar_expect_tx(&full[stage], kStageBytes);
          uint8_t* dst = s_a + stage * kStageBytes;
#pragma unroll
          for (int c = 0; c < CHUNKS; ++c) {
            const __nv_bfloat16* src = (c < 4) ? a.in_a : a.in_b;
            bulk_g2s(dst + c * kRowBytes, src + off + (size_t)(c & 3) * a.Wp * 8, kRowBytes, &full[stage]);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer ---------------------------------------------------------------
    // All 32 lanes run the loop and the barrier waits (warp-uniform control flow: descriptors stay
    // in uniform registers and UTCHMMA issues back to back); one elected lane issues MMAs/commits.
    mbar_wait(wbar, 0);
    const uint64_t a_desc0 = smem_desc(smem_u32(s_a), FIRST ? 16 : kChunkBytes, 128);
    const uint64_t b_desc0 = smem_desc(smem_u32(s_w), kBPlane, 128);
    // Ring positions are running counters over the CTA's whole life: input row #irow sits in stage
    // irow % kStages (phase irow / kStages), output row #o accumulates in slot o % kSlots (use o / kSlots).
    uint32_t irow = 0, obase = 0;
    // cnt output rows starting at slot s0; dyslot of the first: 0 = dy+1 (output above), 1 = dy 0, 2 = dy-1
    auto issue = [&](uint64_t a_stage, uint32_t s0, int cnt, int dyslot) {
      const uint32_t idesc = idesc_bf16_m128(32 * cnt);
      const uint32_t d = tmem + s0 * 32;
      const uint64_t b_row = b_desc0 + (uint64_t)(dyslot * (512 >> 4));
      if constexpr (FIRST) {
        mma_bf16_ss(d, a_stage, b_row, idesc, 1u);                                // dx = 0, 1
        mma_bf16_ss(d, a_stage + 2, b_row + (uint64_t)(kBTile >> 4), idesc, 1u);  // dx = 2, (zero)
      } else {
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
#pragma unroll
          for (int kk = 0; kk < CHUNKS / 2; ++kk) {
            const uint64_t ad = a_stage + (uint64_t)((kk * 2 * kChunkBytes + dx * 16) >> 4);
            const uint64_t bd = b_row + (uint64_t)(((dx * (CHUNKS / 2) + kk) * kBTile) >> 4);
            mma_bf16_ss(d, ad, bd, idesc, 1u);
          }
        }
      }
    };
    // One input row, general form (unit borders, ring wrap).  r = relative output row centred on it.
    auto row_general = [&](int r, int rows, bool first, bool last) {
      const int lo = max(r - 1, 0), hi = min(r + 1, rows - 1);  // relative output rows it contributes to
      // slots touched for the first time must have been drained and re-initialised with the bias
      for (int o = first ? lo : r + 1; o <= hi; ++o) {
        const uint32_t og = obase + o;
        mbar_wait(&acc_free[og & (kSlots - 1)], (og / kSlots) & 1);
      }
      const uint32_t stage = irow & (kStages - 1);
      mbar_wait(&full[stage], (irow / kStages) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t a_stage = a_desc0 + (uint64_t)(stage * (kStageBytes >> 4));
        const uint32_t s0 = (obase + lo) & (kSlots - 1);
        const int cnt = hi - lo + 1, c1 = min(cnt, kSlots - (int)s0), dyslot = lo - (r - 1);
        issue(a_stage, s0, c1, dyslot);
        if (c1 < cnt) issue(a_stage, 0, cnt - c1, dyslot + c1);  // the slot ring wrapped
        mma_commit(&empty[stage]);
        // output row r-1 has now received its last tap; at the bottom image border so has row r
        if (r >= 1) mma_commit(&acc_ready[(obase + r - 1) & (kSlots - 1)]);
        if (last && r == rows - 1) mma_commit(&acc_ready[(obase + r) & (kSlots - 1)]);
      }
      __syncwarp();
      ++irow;
    };
    for (int u = blockIdx.x; u < a.units; u += gridDim.x) {
      int n, sx, ys, ye;
      decode(u, n, sx, ys, ye);
      const int yi0 = max(ys - 1, 0), yi1 = min(ye, H - 1), rows = ye - ys;
      const int r0 = yi0 - ys, r1 = yi1 - ys;  // relative input rows of the unit: r0 in {-1, 0}, r1 in {rows-1, rows}
      int r = r0;
      for (; r <= min(0, r1); ++r) row_general(r, rows, r == r0, r == r1);
      // interior rows 1 .. rows-2: three output rows r-1..r+1, one new slot (r+1), one finished row (r-1)
      for (; r <= rows - 2; ++r, ++irow) {
        const uint32_t og = obase + r + 1;  // the newly touched output row
        mbar_wait(&acc_free[og & (kSlots - 1)], (og / kSlots) & 1);
        const uint32_t stage = irow & (kStages - 1);
        mbar_wait(&full[stage], (irow / kStages) & 1);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t a_stage = a_desc0 + (uint64_t)(stage * (kStageBytes >> 4));
          const uint32_t s0 = (og - 2) & (kSlots - 1);
          if (s0 <= kSlots - 3) {
            issue(a_stage, s0, 3, 0);
          } else {
            issue(a_stage, s0, kSlots - (int)s0, 0);
            issue(a_stage, 0, 3 - (kSlots - (int)s0), kSlots - (int)s0);
          }
          mma_commit(&empty[stage]);
          mma_commit(&acc_ready[s0]);
        }
        __syncwarp();
      }
      for (; r <= r1; ++r) row_general(r, rows, false, r == r1);
      obase += rows;
    }
  } else {
    // ---------------- epilogue -------------------------------------------------------------------
    const int q = warp & 3, g = (warp - 2) >> 2;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    uint32_t bias_u[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) bias_u[j] = __float_as_uint(__ldg(a.bias + j));
    for (int s = g; s < kSlots; s += 2) tmem_st32(tmem + lane_base + s * 32, bias_u);
    tmem_wait_st();
    tc_fence_before();
    __syncwarp();
    if (lane == 0)
      for (int s = g; s < kSlots; s += 2) mbar_arrive(&acc_free[s]);
    uint32_t obase = 0;  // running output-row counter (same sequence as the issuer's)
    for (int u = blockIdx.x; u < a.units; u += gridDim.x) {
      int n, sx, ys, ye;
      decode(u, n, sx, ys, ye);
      const int x = sx * kTileX + q * 32 + lane;
      const bool x_ok = x < a.W;
      const int rows = ye - ys;
      for (int i = (g ^ obase) & 1; i < rows; i += 2) {  // group g owns the output rows with (counter & 1) == g
        const uint32_t og = obase + i;
        const int y = ys + i, s = og & (kSlots - 1);
        mbar_wait(&acc_ready[s], (og / kSlots) & 1);
        tc_fence_after();
        uint32_t v[32];
        tmem_ld32(tmem + lane_base + s * 32, v);
        tmem_wait_ld();
        tmem_st32(tmem + lane_base + s * 32, bias_u);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc_free[s]);
        if constexpr (!LAST) {
          if (x_ok) {
            uint4* dst = reinterpret_cast<uint4*>(a.out + ((((size_t)n * H + y) * 4) * a.Wp + (x + 1)) * 8);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint4 o;
              o.x = pack_bf16_relu(v[8 * c + 0], v[8 * c + 1]);
              o.y = pack_bf16_relu(v[8 * c + 2], v[8 * c + 3]);
              o.z = pack_bf16_relu(v[8 * c + 4], v[8 * c + 5]);
              o.w = pack_bf16_relu(v[8 * c + 6], v[8 * c + 7]);
              dst[(size_t)c * a.Wp] = o;
            }
          }
        } else {
          if (x_ok) {
            const long long bx = n * a.vx.sn + (long long)y * a.vx.sy + (long long)x * a.vx.sx;
            float p[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) p[c] = __ldg(a.x + bx + c * a.vx.sc);
            float A[24];
#pragma unroll
            for (int j = 0; j < 24; ++j) A[j] = fast_tanh(__uint_as_float(v[j]));
#pragma unroll
            for (int it = 0; it < 8; ++it)
              if (it < a.iters) {
#pragma unroll
                for (int c = 0; c < 3; ++c) p[c] = fmaf(a.sign * A[3 * it + c] * p[c], 1.f - p[c], p[c]);
              }
#pragma unroll
            for (int c = 0; c < 3; ++c) a.y[bx + c * a.vx.sc] = p[c];
            if (a.maps) {
              const long long bm = n * a.vm.sn + (long long)y * a.vm.sy + (long long)x * a.vm.sx;
#pragma unroll
              for (int j = 0; j < 24; ++j) a.maps[bm + j * a.vm.sc] = A[j];
            }
          }
        }
      }
      obase += rows;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// Input staging for conv1: fp32 image (any layout) -> bf16 [n][y][Wp][8] (RGB + 5 zero channels),
// pixel x at column x+1 (column 0 and the columns past W stay zero from the workspace memset).
// One thread per pixel: 12 B read, one 16 B store.
__global__ void __launch_bounds__(256)
dce_pack_input_kernel(const float* __restrict__ x, ImgView v, __nv_bfloat16* __restrict__ out, int H, int W, int Wp) {
  const int xx = blockIdx.x * blockDim.x + threadIdx.x, yy = blockIdx.y;
  const long long n = blockIdx.z;
  if (xx >= W) return;
  const float* p = x + n * v.sn + (long long)yy * v.sy + (long long)xx * v.sx;
  const __nv_bfloat162 rg = __floats2bfloat162_rn(__ldg(p), __ldg(p + v.sc));
  const __nv_bfloat162 b0 = __floats2bfloat162_rn(__ldg(p + 2 * v.sc), 0.f);
  uint4 o;
  o.x = *reinterpret_cast<const uint32_t*>(&rg);
  o.y = *reinterpret_cast<const uint32_t*>(&b0);
  o.z = 0u; o.w = 0u;
  *reinterpret_cast<uint4*>(out + ((n * H + yy) * (long long)Wp + (xx + 1)) * 8) = o;
}

// Segments per strip: minimise waves x (rows per segment + halo + pipeline fill) over the grid.
int choose_segments(int n, int strips, int H, int ctas) {
  long long best_cost = -1;
  int best = 1;
  for (int segs = 1; segs <= 64 && segs * 8 <= H + 7; ++segs) {
    long long units = (long long)n * strips * segs;
    long long waves = (units + ctas - 1) / ctas;
    int rows = (H + segs - 1) / segs;
    long long cost = waves * (rows + 2 + 6);
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = segs; }
  }
  return best;
}

template <int CHUNKS, bool LAST, bool FIRST = false>
int launch_tc(llie_ctx* ctx, cudaStream_t st, ConvArgs a) {
  constexpr size_t smem = (FIRST ? 2 * kBTile : 3 * (CHUNKS / 2) * kBTile) +
                          (size_t)kStages * CHUNKS * (FIRST ? kFirstRowBytes : kChunkBytes) +
                          (2 * kStages + 2 * kSlots + 1) * sizeof(uint64_t) + 16;
  LLIE_CUDA_OK(cudaFuncSetAttribute(dce_conv_tc_kernel<CHUNKS, LAST, FIRST>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)smem));
  const int ctas = ctx->sm_count;
  a.segs = choose_segments(a.N, a.strips, a.H, ctas);
  a.seg_rows = (a.H + a.segs - 1) / a.segs;
  a.segs = (a.H + a.seg_rows - 1) / a.seg_rows;
  a.units = a.N * a.strips * a.segs;
  const int grid = a.units < ctas ? a.units : ctas;
  LLIE_LAUNCH(ctx, st,
              FIRST ? "dce_conv_tc_first_kernel"
                    : LAST ? "dce_conv_tc_last_kernel" : (CHUNKS == 4 ? "dce_conv_tc_k32_kernel" : "dce_conv_tc_k64_kernel"),
              (dce_conv_tc_kernel<CHUNKS, LAST, FIRST><<<grid, kThreads, smem, st>>>(a)));
  return LLIE_OK;
}

}  // namespace

// Weight tiles.  Layers l = 1..6: for every (dx, k16) a 96 x 16 bf16 tile, row n = dyslot*32 + cout
// with dyslot 0/1/2 <-> kernel row ky = 2/1/0, stored as two 8-channel planes of 96 x 16 B (K-major,
// no swizzle).  Layer 0 (conv1): two tiles whose K halves are horizontal taps -- tile 0 = (dx=0 |
// dx=1), tile 1 = (dx=2 | zeros) -- each half holding the 3 input channels + 5 zero channels.
struct DceTcPack {
  size_t tile_off[7];  // element offsets into d_bf16 (layers 0..6)
  float* d_bias = nullptr;  // 7 x 32
  // geometry of the activation workspace of the last forward (for the debug hook / pad zeroing)
  int n = 0, h = 0, w = 0, wp = 0;
  __nv_bfloat16* xin = nullptr;  // packed input [n][h][wp][8]
  __nv_bfloat16* act[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};
static std::map<llie_ctx*, DceTcPack> g_packs;  // keyed by ctx (ctx is single-threaded by contract)

int dce_pack_bf16(llie_ctx* ctx, DceWeights* dw) {
  DceTcPack& pk = g_packs[ctx];
  cudaFree(pk.d_bias);
  pk.d_bias = nullptr;
  const float* raw = dw->host_raw.data();
  std::vector<__nv_bfloat16> tiles;
  std::vector<float> biases(7 * 32, 0.f);
  size_t src = 0;
  for (int l = 0; l < 7; ++l) {
    const int co = kCout[l], ci = kCin[l];
    const float* Wl = raw + src;
    const float* bl = raw + src + (size_t)co * ci * 9;
    for (int o = 0; o < co; ++o) biases[l * 32 + o] = bl[o];
    pk.tile_off[l] = tiles.size();
    if (l == 0) {
      tiles.resize(tiles.size() + (size_t)2 * 96 * 16, __float2bfloat16(0.f));
      __nv_bfloat16* T = tiles.data() + pk.tile_off[l];
      for (int tile = 0; tile < 2; ++tile)
        for (int kc = 0; kc < 2; ++kc) {
          const int dx = tile * 2 + kc;
          if (dx > 2) continue;
          for (int nrow = 0; nrow < 96; ++nrow)
            for (int e = 0; e < 3; ++e) {
              const int o = nrow & 31, ky = 2 - (nrow >> 5);
              T[(((size_t)tile * 2 + kc) * 96 + nrow) * 8 + e] = __float2bfloat16(Wl[(((size_t)o * ci + e) * 3 + ky) * 3 + dx]);
            }
        }
    } else {
      const int k16 = ci / 16;
      tiles.resize(tiles.size() + (size_t)3 * k16 * 96 * 16, __float2bfloat16(0.f));
      __nv_bfloat16* T = tiles.data() + pk.tile_off[l];
      for (int dx = 0; dx < 3; ++dx)
        for (int kk = 0; kk < k16; ++kk)
          for (int kc = 0; kc < 2; ++kc)
            for (int nrow = 0; nrow < 96; ++nrow)
              for (int e = 0; e < 8; ++e) {
                const int o = nrow & 31, ky = 2 - (nrow >> 5), cin = kk * 16 + kc * 8 + e;
                float val = o < co ? Wl[(((size_t)o * ci + cin) * 3 + ky) * 3 + dx] : 0.f;
                T[(((size_t)(dx * k16 + kk) * 2 + kc) * 96 + nrow) * 8 + e] = __float2bfloat16(val);
              }
    }
    src += (size_t)co * ci * 9 + co;
  }
  LLIE_CUDA_OK(cudaMalloc(&dw->d_bf16, tiles.size() * sizeof(__nv_bfloat16)));
  LLIE_CUDA_OK(cudaMemcpy(dw->d_bf16, tiles.data(), tiles.size() * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice));
  LLIE_CUDA_OK(cudaMalloc(&pk.d_bias, biases.size() * sizeof(float)));
  LLIE_CUDA_OK(cudaMemcpy(pk.d_bias, biases.data(), biases.size() * sizeof(float), cudaMemcpyHostToDevice));
  return LLIE_OK;
}

void dce_tc_release(llie_ctx* ctx) {
  auto it = g_packs.find(ctx);
  if (it == g_packs.end()) return;
  cudaFree(it->second.d_bias);
  g_packs.erase(it);
}

int dce_forward_bf16(llie_ctx* ctx, const float* in, float* out, float* maps, int n, int h, int w, int layout,
                     const llie_dce_params& p, cudaStream_t st) {
  DceWeights* dw = ctx->dce;
  LLIE_REQUIRE(dw != nullptr && dw->d_bf16 != nullptr, LLIE_E_STATE, "dce_forward: call llie_dce_set_weights first");
  DceTcPack& pk = g_packs[ctx];
  const int strips = (w + kTileX - 1) / kTileX, wp = strips * kTileX + 2;
  const size_t act_elems_per_img = (size_t)h * 4 * wp * 8, in_elems_per_img = (size_t)h * wp * 8;
  // frames per pass: six bf16 activations + the packed input per frame (0.93 GB at 1080p), kept under
  // ~24 GiB, and the passes balanced (64 frames -> 22 + 21 + 21, not 25 + 25 + 14)
  long long chunk = (24ll << 30) / (long long)((6 * act_elems_per_img + in_elems_per_img) * 2);
  chunk = chunk < 1 ? 1 : (chunk > n ? n : chunk);
  const long long passes = (n + chunk - 1) / chunk;
  chunk = (n + passes - 1) / passes;
  void* ws;
  // + slack: conv1's staged rows run 2 pixels past a row's end (zero-weighted, but must be readable)
  const size_t ws_bytes = (size_t)chunk * (6 * act_elems_per_img + in_elems_per_img) * 2 + 256;
  LLIE_TRY(ctx->workspace("dce.bf16.act", ws_bytes, &ws));
  if (pk.n != (int)chunk || pk.h != h || pk.w != w || pk.act[0] != (__nv_bfloat16*)ws) {
    // pad columns must read as zero and move with the geometry: clear on every geometry change
    LLIE_CUDA_OK(cudaMemsetAsync(ws, 0, ws_bytes, st));
    pk.n = (int)chunk; pk.h = h; pk.w = w; pk.wp = wp;
    for (int i = 0; i < 6; ++i) pk.act[i] = (__nv_bfloat16*)ws + (size_t)i * chunk * act_elems_per_img;
    pk.xin = (__nv_bfloat16*)ws + (size_t)6 * chunk * act_elems_per_img;
  }
  const ImgView vx = make_view(layout, 3, h, w), vm = make_view(layout, 24, h, w);
  for (long long n0 = 0; n0 < n; n0 += chunk) {
    const int nb = (int)((n - n0 < chunk) ? (n - n0) : chunk);
    const float* x = in + n0 * (long long)h * w * 3;
    {
      dim3 grid((w + 255) / 256, h, nb);
      LLIE_LAUNCH(ctx, st, "dce_pack_input_kernel",
                  dce_pack_input_kernel<<<grid, 256, 0, st>>>(x, vx, pk.xin, h, w, wp));
    }
    ConvArgs a;
    memset(&a, 0, sizeof(a));
    a.N = nb; a.H = h; a.W = w; a.Wp = wp; a.strips = strips;
    auto layer = [&](int l, const __nv_bfloat16* ia, const __nv_bfloat16* ib, __nv_bfloat16* o) {
      a.in_a = ia; a.in_b = ib; a.out = o;
      a.wpk = dw->d_bf16 + pk.tile_off[l];
      a.bias = pk.d_bias + l * 32;
    };
    layer(0, pk.xin, nullptr, pk.act[0]); LLIE_TRY((launch_tc<1, false, true>(ctx, st, a)));
    layer(1, pk.act[0], nullptr, pk.act[1]); LLIE_TRY((launch_tc<4, false>(ctx, st, a)));
    layer(2, pk.act[1], nullptr, pk.act[2]); LLIE_TRY((launch_tc<4, false>(ctx, st, a)));
    layer(3, pk.act[2], nullptr, pk.act[3]); LLIE_TRY((launch_tc<4, false>(ctx, st, a)));
    layer(4, pk.act[2], pk.act[3], pk.act[4]); LLIE_TRY((launch_tc<8, false>(ctx, st, a)));
    layer(5, pk.act[1], pk.act[4], pk.act[5]); LLIE_TRY((launch_tc<8, false>(ctx, st, a)));
    layer(6, pk.act[0], pk.act[5], nullptr);
    a.x = x; a.y = out + n0 * (long long)h * w * 3;
    a.maps = maps ? maps + n0 * (long long)h * w * 24 : nullptr;
    a.vx = vx; a.vm = vm; a.iters = p.iters; a.sign = p.curve_sign >= 0 ? 1.f : -1.f;
    LLIE_TRY((launch_tc<8, true>(ctx, st, a)));
  }
  return LLIE_OK;
}

// Test hook: activation `which` (1..6) of frame `frame` from the last bf16 forward, as fp32 [32][h][w].
int dce_debug_activation(llie_ctx* ctx, int which, int frame, float* host_out) {
  auto it = g_packs.find(ctx);
  LLIE_REQUIRE(it != g_packs.end() && it->second.act[0], LLIE_E_STATE, "no bf16 forward has run on this ctx");
  DceTcPack& pk = it->second;
  LLIE_REQUIRE(which >= 1 && which <= 6 && frame >= 0 && frame < pk.n && host_out, LLIE_E_ARG, "bad activation query");
  const size_t per = (size_t)pk.h * 4 * pk.wp * 8;
  std::vector<__nv_bfloat16> buf(per);
  LLIE_CUDA_OK(cudaDeviceSynchronize());
  LLIE_CUDA_OK(cudaMemcpy(buf.data(), pk.act[which - 1] + (size_t)frame * per, per * 2, cudaMemcpyDeviceToHost));
  for (int y = 0; y < pk.h; ++y)
    for (int c = 0; c < 32; ++c)
      for (int x = 0; x < pk.w; ++x)
        host_out[((size_t)c * pk.h + y) * pk.w + x] =
            __bfloat162float(buf[(((size_t)y * 4 + (c >> 3)) * pk.wp + (x + 1)) * 8 + (c & 7)]);
  return LLIE_OK;
}

}  // namespace llie
