// Zero-DCE, bf16 tensor-core path (config 2): host side.  All seven DCE-Net layers run on the
// 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM) through the fused chain kernel of
// dce_chain.cuh, as three launches per pass: conv1+conv2+conv3, conv4+conv5, conv6+conv7 (tanh and
// the 8 curve iterations fused into conv7's epilogue).  x4 and x6 never reach HBM; x1, x2, x3, x5
// are written once and read once.  (north_star kernels (2) and (3))
// No reference file exists to cite (/root/reference/README.md:1-2).
//
// Activations live in HBM as [n][y][chunk=C/8][Wp][8 ch] bf16, image pixel x at column x + kPadL,
// every other column zero, so one input row of a strip is C/8 contiguous runs of 132 pixels x 16 B
// that a 1-D bulk async copy (TMA engine) drops into shared memory in the K-major no-swizzle layout
// tcgen05 wants; a horizontal tap dx is the A descriptor's start address + 16*dx bytes.
#include "dce_chain.cuh"
#include "dce_weights.cuh"

#include <cstring>

#ifdef LLIE_MBAR_TRACE
extern "C" int llie_debug_set_trace(void* host_mapped) {
  return (int)cudaMemcpyToSymbol(g_llie_trace, &host_mapped, sizeof(void*));
}
#endif

namespace llie {

namespace {
using namespace chain;

// Input staging for conv1: fp32 image (any layout) -> bf16 [n][y][Wp][8] (RGB + 5 zero channels),
// pixel x at column x + kPadL (all other columns stay zero from the workspace memset).
// One thread per pixel: 12 B read, one 16 B store.
// kSplit: channels = [r_hi g_hi b_hi r_lo g_lo b_lo 0 0] with hi = bf16(v), lo = bf16(v - hi).
template <bool kSplit, typename TIn>
__global__ void __launch_bounds__(256)
dce_pack_input_kernel(const TIn* __restrict__ x, ImgView v, __nv_bfloat16* __restrict__ out, int H, int W, int Wp) {
  const int xx = blockIdx.x * blockDim.x + threadIdx.x, yy = blockIdx.y;
  const long long n = blockIdx.z;
  if (xx >= W) return;
  const TIn* p = x + n * v.sn + (long long)yy * v.sy + (long long)xx * v.sx;
  // an integer b means b / 255 (65535); float and bfloat16 samples are the value itself
  constexpr float kIn = std::is_same<TIn, uint8_t>::value ? 1.f / 255.f : (std::is_same<TIn, uint16_t>::value ? 1.f / 65535.f : 1.f);
  const float r = kIn * (float)__ldg(p), g = kIn * (float)__ldg(p + v.sc), b = kIn * (float)__ldg(p + 2 * v.sc);
  const __nv_bfloat162 rg = __floats2bfloat162_rn(r, g);
  uint4 o;
  o.x = *reinterpret_cast<const uint32_t*>(&rg);
  if constexpr (kSplit) {
    const __nv_bfloat16 bh = __float2bfloat16_rn(b);
    const __nv_bfloat162 b_rl = __halves2bfloat162(bh, __float2bfloat16_rn(r - __low2float(rg)));
    const __nv_bfloat162 gb_l = __floats2bfloat162_rn(g - __high2float(rg), b - __bfloat162float(bh));
    o.y = *reinterpret_cast<const uint32_t*>(&b_rl);
    o.z = *reinterpret_cast<const uint32_t*>(&gb_l);
    o.w = 0u;
  } else {
    const __nv_bfloat162 b0 = __floats2bfloat162_rn(b, 0.f);
    o.y = *reinterpret_cast<const uint32_t*>(&b0);
    o.z = 0u; o.w = 0u;
  }
  *reinterpret_cast<uint4*>(out + ((n * H + yy) * (long long)Wp + (xx + kPadL)) * 8) = o;
}

// Segments per strip: minimise waves x (rows per segment + halo + pipeline fill) over the grid.
int choose_segments(int n, int strips, int H, int ctas, int extra_rows) {
  long long best_cost = -1;
  int best = 1;
  for (int segs = 1; segs <= 64 && segs * 8 <= H + 7; ++segs) {
    long long units = (long long)n * strips * segs;
    long long waves = (units + ctas - 1) / ctas;
    int rows = (H + segs - 1) / segs;
    long long cost = waves * (rows + extra_rows);
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = segs; }
  }
  return best;
}

template <int GID>
int launch_chain(llie_ctx* ctx, cudaStream_t st, ChainArgs a, const char* name) {
  constexpr Group G = group_cfg(GID);
  constexpr int smem = smem_bytes(G);
  LLIE_CUDA_OK(cudaFuncSetAttribute(dce_chain_kernel<GID>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int ctas = ctx->sm_count;
  a.strips = (a.W + valid_px(G) - 1) / valid_px(G);
  a.segs = choose_segments(a.N, a.strips, a.H, ctas, 8 + (2 + G.lag) * (G.S - 1));
  a.seg_rows = (a.H + a.segs - 1) / a.segs;
  a.segs = (a.H + a.seg_rows - 1) / a.seg_rows;
  a.units = a.N * a.strips * a.segs;
  const int grid = a.units < ctas ? a.units : ctas;
  LLIE_LAUNCH(ctx, st, name, (dce_chain_kernel<GID><<<grid, n_threads(G), smem, st>>>(a)));
  return LLIE_OK;
}

}  // namespace

// Weight tiles.  Layers l = 1..6: for every (dx, k16) a 96 x 16 bf16 tile, row n = dyslot*32 + cout
// with dyslot 0/1/2 <-> kernel row ky = 2/1/0, stored as two 8-channel planes of 96 x 16 B (K-major,
// no swizzle).  Layer 0 (conv1): two tiles whose K halves are horizontal taps -- tile 0 = (dx=0 |
// dx=1), tile 1 = (dx=2 | zeros) -- each half holding the 3 input channels + 5 zero channels.
struct DceTcPack {
  size_t tile_off[7];        // element offsets into d_bf16 (layers 0..6)
  size_t tile_off_split[7];  // ... of the hi/lo weight tiles of the split (fp32-accurate) path
  bool split = false;        // layout of the activations of the last forward
  float h_bias[7 * 32] = {0};  // per layer, host copy (passed to the kernels by value)
  // geometry of the activation workspace of the last forward (for the debug hook / pad zeroing)
  int n = 0, h = 0, w = 0, wp = 0;
  __nv_bfloat16* xin = nullptr;  // packed input [n][h][wp][8]
  bool keep_all = false;         // test hook: also write x4 / x6 (never needed by a later launch) to HBM
  bool unfused = false;          // test hook: one launch per layer (the same kernel with S = 1)
  __nv_bfloat16* act[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};
// Owned by the ctx (llie_ctx::tc), created on first use: no process-wide state, so host threads that each drive
// their own ctx never touch a shared container.
static DceTcPack& tc_pack(llie_ctx* ctx) {
  if (!ctx->tc) ctx->tc = new DceTcPack;
  return *ctx->tc;
}

int dce_pack_bf16(llie_ctx* ctx, DceWeights* dw) {
  DceTcPack& pk = tc_pack(ctx);
  const float* raw = dw->host_raw.data();
  std::vector<__nv_bfloat16> tiles;
  float* biases = pk.h_bias;
  memset(biases, 0, sizeof(pk.h_bias));
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
  // Split path: every weight as hi = bf16(w) and lo = bf16(w - hi).  Layers 1..6: the 3*k16 tiles of w_hi, then
  // those of w_lo.  Layer 0: the input chunk is [rgb_hi rgb_lo 0 0], tiles 0/1 = w_hi under the hi AND the lo
  // channels (a_hi*w_hi + a_lo*w_hi), tiles 2/3 = w_lo under the hi channels (a_hi*w_lo).
  src = 0;
  auto hi_of = [](float v) { return __float2bfloat16(v); };
  auto lo_of = [](float v) { return __float2bfloat16(v - __bfloat162float(__float2bfloat16(v))); };
  for (int l = 0; l < 7; ++l) {
    const int co = kCout[l], ci = kCin[l];
    const float* Wl = raw + src;
    pk.tile_off_split[l] = tiles.size();
    if (l == 0) {
      tiles.resize(tiles.size() + (size_t)4 * 96 * 16, __float2bfloat16(0.f));
      __nv_bfloat16* T = tiles.data() + pk.tile_off_split[l];
      for (int part = 0; part < 2; ++part)
        for (int tile = 0; tile < 2; ++tile)
          for (int kc = 0; kc < 2; ++kc) {
            const int dx = tile * 2 + kc;
            if (dx > 2) continue;
            for (int nrow = 0; nrow < 96; ++nrow)
              for (int e = 0; e < 3; ++e) {
                const int o = nrow & 31, ky = 2 - (nrow >> 5);
                const float wv = Wl[(((size_t)o * ci + e) * 3 + ky) * 3 + dx];
                __nv_bfloat16* row = T + (((size_t)(part * 2 + tile) * 2 + kc) * 96 + nrow) * 8;
                if (part == 0) { row[e] = hi_of(wv); row[3 + e] = hi_of(wv); }
                else row[e] = lo_of(wv);
              }
          }
    } else {
      const int k16 = ci / 16;
      tiles.resize(tiles.size() + (size_t)2 * 3 * k16 * 96 * 16, __float2bfloat16(0.f));
      __nv_bfloat16* T = tiles.data() + pk.tile_off_split[l];
      for (int part = 0; part < 2; ++part)
        for (int dx = 0; dx < 3; ++dx)
          for (int kk = 0; kk < k16; ++kk)
            for (int kc = 0; kc < 2; ++kc)
              for (int nrow = 0; nrow < 96; ++nrow)
                for (int e = 0; e < 8; ++e) {
                  const int o = nrow & 31, ky = 2 - (nrow >> 5), cin = kk * 16 + kc * 8 + e;
                  const float wv = o < co ? Wl[(((size_t)o * ci + cin) * 3 + ky) * 3 + dx] : 0.f;
                  T[((((size_t)part * 3 * k16 + dx * k16 + kk) * 2 + kc) * 96 + nrow) * 8 + e] = part ? lo_of(wv) : hi_of(wv);
                }
    }
    src += (size_t)co * ci * 9 + co;
  }
  LLIE_CUDA_OK(cudaMalloc(&dw->d_bf16, tiles.size() * sizeof(__nv_bfloat16)));
  LLIE_CUDA_OK(cudaMemcpy(dw->d_bf16, tiles.data(), tiles.size() * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice));
  return LLIE_OK;
}

void dce_tc_release(llie_ctx* ctx) {
  delete ctx->tc;
  ctx->tc = nullptr;
}

// split = false: bf16 operands (config 2).  split = true: hi/lo bf16 pairs, three MMAs per product -- fp32-accurate
// results (config 1) from the same tensor-core kernel, one launch per layer.
static int dce_forward_tc(llie_ctx* ctx, const void* in_v, void* out_v, float* maps, int n, int h, int w, int layout,
                          const llie_dce_params& p, cudaStream_t st, bool split, int iot = kIoF32) {
  const size_t esz = io_bytes(iot);  // bytes per image sample
  DceWeights* dw = ctx->dce;
  LLIE_REQUIRE(dw != nullptr && dw->d_bf16 != nullptr, LLIE_E_STATE, "dce_forward: call llie_dce_set_weights first");
  DceTcPack& pk = tc_pack(ctx);
  // row pitch in pixels: kPadL zero columns on the left, and enough zero columns on the right for the
  // last strip's 132-pixel staged rows (a strip's first staged column is < W + kPadL)
  const int wp = w + kPadL + kRowPx + 4;
  const int planes = split ? 8 : 4;  // 8-channel chunk planes per activation row
  const size_t act_elems_per_img = (size_t)h * planes * wp * 8, in_elems_per_img = (size_t)h * wp * 8;
  // frames per pass: six bf16 activations + the packed input per frame (about 1 GB at 1080p), kept
  // under ~24 GiB, and the passes balanced (64 frames -> 22 + 21 + 21, not 25 + 25 + 14)
  long long chunk = (24ll << 30) / (long long)((6 * act_elems_per_img + in_elems_per_img) * 2);
  chunk = chunk < 1 ? 1 : (chunk > n ? n : chunk);
  const long long passes = (n + chunk - 1) / chunk;
  chunk = (n + passes - 1) / passes;
  void* ws;
  const size_t ws_bytes = (size_t)chunk * (6 * act_elems_per_img + in_elems_per_img) * 2 + 256;
  LLIE_TRY(ctx->workspace("dce.bf16.act", ws_bytes, &ws));
  if (pk.n != (int)chunk || pk.h != h || pk.w != w || pk.split != split || pk.act[0] != (__nv_bfloat16*)ws) {
    // pad columns must read as zero and move with the geometry: clear on every geometry change
    LLIE_CUDA_OK(cudaMemsetAsync(ws, 0, ws_bytes, st));
    pk.n = (int)chunk; pk.h = h; pk.w = w; pk.wp = wp; pk.split = split;
    for (int i = 0; i < 6; ++i) pk.act[i] = (__nv_bfloat16*)ws + (size_t)i * chunk * act_elems_per_img;
    pk.xin = (__nv_bfloat16*)ws + (size_t)6 * chunk * act_elems_per_img;
  }
  const ImgView vx = make_view(layout, 3, h, w), vm = make_view(layout, 24, h, w);
  // Several passes (64 frames = 22 + 21 + 21): the packed input is only read by the first chain of a pass, so the
  // NEXT pass's pack kernel (HBM-bound, no shared memory: it fits beside the persistent chain CTAs) runs on the ctx's
  // second stream underneath conv4-5 / conv6-7 of the current pass.
  const bool overlap_pack = !split && !pk.unfused && n > chunk;
  if (overlap_pack) LLIE_TRY(ensure_aux_stream(ctx));
  auto pack = [&](long long n0, cudaStream_t s2) -> int {
    const int nb = (int)((n - n0 < chunk) ? (n - n0) : chunk);
    const float* x = reinterpret_cast<const float*>(static_cast<const char*>(in_v) + (size_t)n0 * h * w * 3 * esz);
    dim3 grid((w + 255) / 256, h, nb);
    if (split)
      LLIE_LAUNCH(ctx, s2, "dce_pack_input_split_kernel",
                  (dce_pack_input_kernel<true, float><<<grid, 256, 0, s2>>>(x, vx, pk.xin, h, w, wp)));
    else if (iot == kIoU8)
      LLIE_LAUNCH(ctx, s2, "dce_pack_input_kernel",
                  (dce_pack_input_kernel<false, uint8_t><<<grid, 256, 0, s2>>>(reinterpret_cast<const uint8_t*>(x), vx, pk.xin, h, w, wp)));
    else if (iot == kIoU16)
      LLIE_LAUNCH(ctx, s2, "dce_pack_input_kernel",
                  (dce_pack_input_kernel<false, uint16_t><<<grid, 256, 0, s2>>>(reinterpret_cast<const uint16_t*>(x), vx, pk.xin, h, w, wp)));
    else if (iot == kIoBF16)
      LLIE_LAUNCH(ctx, s2, "dce_pack_input_kernel",
                  (dce_pack_input_kernel<false, __nv_bfloat16><<<grid, 256, 0, s2>>>(reinterpret_cast<const __nv_bfloat16*>(x), vx, pk.xin, h, w, wp)));
    else
      LLIE_LAUNCH(ctx, s2, "dce_pack_input_kernel",
                  (dce_pack_input_kernel<false, float><<<grid, 256, 0, s2>>>(x, vx, pk.xin, h, w, wp)));
    return LLIE_OK;
  };
  for (long long n0 = 0; n0 < n; n0 += chunk) {
    const int nb = (int)((n - n0 < chunk) ? (n - n0) : chunk);
    const float* x = reinterpret_cast<const float*>(static_cast<const char*>(in_v) + (size_t)n0 * h * w * 3 * esz);
    if (!overlap_pack || n0 == 0) LLIE_TRY(pack(n0, st));
    else LLIE_CUDA_OK(cudaStreamWaitEvent(st, ctx->aux_ev[1], 0));  // this pass's input was packed under the previous pass
    ChainArgs a;
    memset(&a, 0, sizeof(a));
    a.N = nb; a.H = h; a.W = w; a.Wp = wp;
    a.x = x; a.y = reinterpret_cast<float*>(static_cast<char*>(out_v) + (size_t)n0 * h * w * 3 * esz);
    a.u8 = iot;  // kIoF32 / kIoU8 / kIoU16 / kIoBF16 = 0 / 1 / 2 / 3
    a.maps = maps ? maps + n0 * (long long)h * w * 24 : nullptr;
    a.vx = vx; a.vm = vm; a.iters = p.iters; a.sign = p.curve_sign >= 0 ? 1.f : -1.f;
    // stage slot <- layer l (0..6): weights, bias, HBM destination (nullptr = keep on chip)
    auto stage = [&](int slot, int l, __nv_bfloat16* o) {
      a.wpk[slot] = dw->d_bf16 + (split ? pk.tile_off_split[l] : pk.tile_off[l]);
      memcpy(a.bias[slot], pk.h_bias + l * 32, 32 * sizeof(float));
      a.out[slot] = o;
    };
    __nv_bfloat16 *x1 = pk.act[0], *x2 = pk.act[1], *x3 = pk.act[2], *x4 = pk.act[3], *x5 = pk.act[4], *x6 = pk.act[5];
    if (split) {
      auto single = [&](int l, const __nv_bfloat16* ia, const __nv_bfloat16* ib, __nv_bfloat16* o) {
        memset(a.ext, 0, sizeof(a.ext));
        stage(0, l, o);
        a.ext[0][0][0] = ia; a.ext[0][0][1] = ib;
      };
      single(0, pk.xin, nullptr, x1); LLIE_TRY((launch_chain<7>(ctx, st, a, "dce_split_first_kernel")));
      single(1, x1, nullptr, x2); LLIE_TRY((launch_chain<8>(ctx, st, a, "dce_split_k32_kernel")));
      single(2, x2, nullptr, x3); LLIE_TRY((launch_chain<8>(ctx, st, a, "dce_split_k32_kernel")));
      single(3, x3, nullptr, x4); LLIE_TRY((launch_chain<8>(ctx, st, a, "dce_split_k32_kernel")));
      single(4, x3, x4, x5); LLIE_TRY((launch_chain<9>(ctx, st, a, "dce_split_k64_kernel")));
      single(5, x2, x5, x6); LLIE_TRY((launch_chain<9>(ctx, st, a, "dce_split_k64_kernel")));
      single(6, x1, x6, nullptr); LLIE_TRY((launch_chain<10>(ctx, st, a, "dce_split_last_kernel")));
    } else if (!pk.unfused) {
      // conv1 + conv2 + conv3: reads the packed image, writes x1, x2, x3
      stage(0, 0, x1); stage(1, 1, x2); stage(2, 2, x3);
      a.ext[0][0][0] = pk.xin;
      LLIE_TRY((launch_chain<0>(ctx, st, a, "dce_chain_c123_kernel")));
      if (overlap_pack && n0 + chunk < n) {  // the packed input is free again: pack the next pass beside conv4..7
        LLIE_CUDA_OK(cudaEventRecord(ctx->aux_ev[0], st));
        LLIE_CUDA_OK(cudaStreamWaitEvent(ctx->aux_stream, ctx->aux_ev[0], 0));
        LLIE_TRY(pack(n0 + chunk, ctx->aux_stream));
        LLIE_CUDA_OK(cudaEventRecord(ctx->aux_ev[1], ctx->aux_stream));
      }
      // conv4 + conv5: reads x3 (twice: as conv4's input and as the first half of conv5's), writes x5
      memset(a.ext, 0, sizeof(a.ext));
      stage(0, 3, pk.keep_all ? x4 : nullptr); stage(1, 4, x5);
      a.ext[0][0][0] = x3; a.ext[1][0][0] = x3;
      LLIE_TRY((launch_chain<1>(ctx, st, a, "dce_chain_c45_kernel")));
      // conv6 + conv7 + tanh + curves: reads x2 | x5, x1 and the image, writes the enhanced image
      memset(a.ext, 0, sizeof(a.ext));
      stage(0, 5, pk.keep_all ? x6 : nullptr); stage(1, 6, nullptr);
      a.ext[0][0][0] = x2; a.ext[0][0][1] = x5; a.ext[1][0][0] = x1;
      LLIE_TRY((launch_chain<2>(ctx, st, a, "dce_chain_c67_kernel")));
    } else {
      auto single = [&](int l, const __nv_bfloat16* ia, const __nv_bfloat16* ib, __nv_bfloat16* o) {
        memset(a.ext, 0, sizeof(a.ext));
        stage(0, l, o);
        a.ext[0][0][0] = ia; a.ext[0][0][1] = ib;
      };
      single(0, pk.xin, nullptr, x1); LLIE_TRY((launch_chain<3>(ctx, st, a, "dce_conv_tc_first_kernel")));
      single(1, x1, nullptr, x2); LLIE_TRY((launch_chain<4>(ctx, st, a, "dce_conv_tc_k32_kernel")));
      single(2, x2, nullptr, x3); LLIE_TRY((launch_chain<4>(ctx, st, a, "dce_conv_tc_k32_kernel")));
      single(3, x3, nullptr, x4); LLIE_TRY((launch_chain<11>(ctx, st, a, "dce_conv_tc_k32_kernel")));  // group 11 = group 4 unless built with -DLLIE_OVF=1 (overflow slots)
      single(4, x3, x4, x5); LLIE_TRY((launch_chain<5>(ctx, st, a, "dce_conv_tc_k64_kernel")));
      single(5, x2, x5, x6); LLIE_TRY((launch_chain<5>(ctx, st, a, "dce_conv_tc_k64_kernel")));
      single(6, x1, x6, nullptr); LLIE_TRY((launch_chain<6>(ctx, st, a, "dce_conv_tc_last_kernel")));
    }
  }
  return LLIE_OK;
}

int dce_forward_bf16(llie_ctx* ctx, const float* in, float* out, float* maps, int n, int h, int w, int layout,
                     const llie_dce_params& p, cudaStream_t st) {
  return dce_forward_tc(ctx, in, out, maps, n, h, w, layout, p, st, false);
}
int dce_forward_int(llie_ctx* ctx, const void* in, void* out, int iot, float* maps, int n, int h, int w, int layout,
                    const llie_dce_params& p, cudaStream_t st) {
  return dce_forward_tc(ctx, in, out, maps, n, h, w, layout, p, st, false, iot);
}
int dce_forward_split(llie_ctx* ctx, const float* in, float* out, float* maps, int n, int h, int w, int layout,
                      const llie_dce_params& p, cudaStream_t st) {
  return dce_forward_tc(ctx, in, out, maps, n, h, w, layout, p, st, true);
}

// Test hooks: mode bit 0 = also store x4 / x6; bit 1 = run one launch per layer.
int dce_debug_mode(llie_ctx* ctx, int mode) {
  DceTcPack& pk = tc_pack(ctx);
  pk.keep_all = (mode & 1) != 0;
  pk.unfused = (mode & 2) != 0;
  return LLIE_OK;
}

// Test hook: activation `which` (1..6) of frame `frame` from the last bf16 forward, as fp32 [32][h][w].
int dce_debug_activation(llie_ctx* ctx, int which, int frame, float* host_out) {
  LLIE_REQUIRE(ctx->tc && ctx->tc->act[0], LLIE_E_STATE, "no bf16 forward has run on this ctx");
  DceTcPack& pk = *ctx->tc;
  LLIE_REQUIRE(which >= 1 && which <= 6 && frame >= 0 && frame < pk.n && host_out, LLIE_E_ARG, "bad activation query");
  const int planes = pk.split ? 8 : 4;
  const size_t per = (size_t)pk.h * planes * pk.wp * 8;
  std::vector<__nv_bfloat16> buf(per);
  LLIE_CUDA_OK(cudaDeviceSynchronize());
  LLIE_CUDA_OK(cudaMemcpy(buf.data(), pk.act[which - 1] + (size_t)frame * per, per * 2, cudaMemcpyDeviceToHost));
  for (int y = 0; y < pk.h; ++y)
    for (int c = 0; c < 32; ++c)
      for (int x = 0; x < pk.w; ++x)
      {
        float v = __bfloat162float(buf[(((size_t)y * planes + (c >> 3)) * pk.wp + (x + kPadL)) * 8 + (c & 7)]);
        if (pk.split) v += __bfloat162float(buf[(((size_t)y * planes + 4 + (c >> 3)) * pk.wp + (x + kPadL)) * 8 + (c & 7)]);
        host_out[((size_t)c * pk.h + y) * pk.w + x] = v;
      }
  return LLIE_OK;
}

}  // namespace llie
