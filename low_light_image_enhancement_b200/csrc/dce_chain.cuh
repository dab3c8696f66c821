// Zero-DCE bf16: a CHAIN of up to three consecutive DCE-Net layers fused in one persistent kernel
// (north_star kernel (2)+(3); SURVEY.md 7.3-C "keep inter-layer activations out of HBM").
// No reference file exists to cite (/root/reference/README.md:1-2).
//
// Each layer ("stage") is the implicit-GEMM formulation of dce_tc.cu: pixels are the MMA's M (128
// consecutive pixels of one image row), the three vertical taps are stacked in N (N = 96), the TMEM
// accumulators of consecutive output rows sit in consecutive 32-column slots, so one MMA group on an
// input row accumulates into output rows y-1, y, y+1.  An input row is used exactly once per stage.
// Fusion: the epilogue of stage s writes its ReLU'd bf16 output row straight into a shared-memory
// ring in the A-operand layout of stage s+1 (and to HBM only if a later group needs it), so the
// intermediate activation never makes the HBM round trip.  Strips overlap in x: a strip computes
// 128 pixels per stage, each stage starting one pixel further right, and owns 128 - 2(S-1) pixels
// of the last stage's output.  Rows outside the image are skipped, columns outside are written as
// zeros (the convolution's zero padding applies to every layer's INPUT).
//
// Warps: s < S = the MMA issuer of stage s (one issuing thread per stage: the accumulation order,
// hence the result, is deterministic), S..3 = bulk-copy producers (one per external input ring),
// 4.. = two epilogue groups of four warps per stage, alternating output rows (warp & 3 = TMEM lane
// quadrant).  Stages are
// ordered against each other only through the ring / accumulator mbarriers.
#pragma once
#include "llie_common.cuh"
#include "sm100_ptx.cuh"

#include <type_traits>

namespace llie {
namespace chain {
using namespace ptx;

constexpr int kTileX = 128;               // pixels per MMA (M)
constexpr int kRowPx = 132;               // staged pixels per ring row
constexpr int kRowBytes = kRowPx * 16;    // one 8-channel chunk of one ring row
constexpr int kBTile = 96 * 16 * 2;       // weight tile of one (dx, k16): N=96 x K=16
constexpr int kBPlane = 96 * 16;          // bytes between the two K halves of a weight tile
constexpr int kPadL = 4;                  // activation column of image pixel x is x + kPadL
constexpr int kMaxStages = 3;
#ifndef LLIE_OVF
#define LLIE_OVF 0
#endif
constexpr int kOvf = LLIE_OVF;
constexpr int kFirstEpiWarp = 4;           // issuer warps + producer warps <= 4

// One input of a stage: `ext` = loaded from HBM by a producer warp (chunks = 1: the packed network
// input, 4: one activation, 8: two concatenated activations), otherwise the ring the previous
// stage's epilogue writes (4 chunks).
struct Src { int ext, chunks, depth; };
struct Stage { int nsrc; Src src[2]; int first, last; };
struct Group { int S; Stage st[kMaxStages]; int lag, slots; int split; int ovf; };
// split = 1: the fp32-accurate path.  Every activation is stored as TWO bf16 tensors, hi = bf16(v) and
// lo = bf16(v - hi) (8 chunk planes per 32 channels: hi 0..3, lo 4..7), every weight likewise, and a
// product a*w is evaluated as a_hi*w_hi + a_lo*w_hi + a_hi*w_lo (three MMAs, fp32 accumulation in TMEM);
// the dropped terms are O(2^-16) relative per product (measured: 1e-5 max-abs on the output).
//
// ovf = 1 (build with -DLLIE_OVF=1): "overflow slots".  With a plain ring of `slots` accumulator slots, the N = 96 MMA
// of an input row whose three output rows straddle the end of the ring has to be split into two MMAs (measured: 106
// cycles for the N=64 + N=32 pair against 56 for one N=96; 2 rows in every `slots`).  Instead each stage gets
// slots + 2 physical slots [P0 | P1 .. P_slots | P_slots+1]; output row y (ABSOLUTE image row) lives in
// P_{1 + y % slots}, and an MMA centred on input row r always writes the three consecutive physical slots starting at
// r % slots.  The last tap of a row with y % slots == slots-1 therefore lands in P0 and the first tap of a row with
// y % slots == 0 in P_slots+1; the epilogue adds the two partial sums of such rows.  Which rows are split depends only
// on y, so results do not depend on how frames are batched or cut into units.  Slots are keyed by y, so across units
// their order of use is irregular: the issuer and each epilogue group keep one parity bit per slot, and a slot always
// belongs to the same epilogue group (rows go to groups by the parity of y; `slots` is even) -- a group that skipped
// uses of a slot would wait on it with the parity of a phase that has not begun.

// G1 = conv1+conv2+conv3, G2 = conv4+conv5, G3 = conv6+conv7; P* = single layers (same kernel, S=1).
__host__ __device__ constexpr Group group_cfg(int id) {
  switch (id) {
    case 0: return Group{3, {Stage{1, {Src{1, 1, 8}, Src{0, 0, 0}}, 1, 0}, Stage{1, {Src{0, 4, 6}, Src{0, 0, 0}}, 0, 0},
                             Stage{1, {Src{0, 4, 6}, Src{0, 0, 0}}, 0, 0}}, 3, 5};
    case 1: return Group{2, {Stage{1, {Src{1, 4, 6}, Src{0, 0, 0}}, 0, 0}, Stage{2, {Src{1, 4, 4}, Src{0, 4, 6}}, 0, 0},
                             Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0}}, 3, kOvf ? 6 : 8, 0, kOvf};
    case 2: return Group{2, {Stage{1, {Src{1, 8, 4}, Src{0, 0, 0}}, 0, 0}, Stage{2, {Src{1, 4, 4}, Src{0, 4, 5}}, 0, 1},
                             Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0}}, 2, kOvf ? 6 : 8, 0, kOvf};
    case 3: return Group{1, {Stage{1, {Src{1, 1, 8}, Src{0, 0, 0}}, 1, 0}, Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0},
                             Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0}}, 2, 16};  // conv1 alone
    case 4: return Group{1, {Stage{1, {Src{1, 4, 8}, Src{0, 0, 0}}, 0, 0}, Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0},
                             Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0}}, 2, 16};  // a 32 -> 32 layer alone
    case 5: return Group{1, {Stage{1, {Src{1, 8, 8}, Src{0, 0, 0}}, 0, 0}, Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0},
                             Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0}}, 2, kOvf ? 6 : 16, 0, kOvf};  // a 64 -> 32 layer alone (summed as in its chain)
    case 6: return Group{1, {Stage{1, {Src{1, 8, 8}, Src{0, 0, 0}}, 0, 1}, Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0},
                             Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0}}, 2, kOvf ? 6 : 16, 0, kOvf};  // conv7 alone (as fused)
    case 11: return Group{1, {Stage{1, {Src{1, 4, 8}, Src{0, 0, 0}}, 0, 0}, Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0},
                              Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0}}, 2, kOvf ? 6 : 16, 0, kOvf};  // conv4 alone (as fused in conv4+5)
    // ---- split (hi/lo) single layers: chunks count hi AND lo planes
    case 7: return Group{1, {Stage{1, {Src{1, 1, 8}, Src{0, 0, 0}}, 1, 0}, Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0},
                             Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0}}, 2, 16, 1};  // conv1
    case 8: return Group{1, {Stage{1, {Src{1, 8, 8}, Src{0, 0, 0}}, 0, 0}, Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0},
                             Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0}}, 2, 16, 1};  // 32 -> 32
    case 9: return Group{1, {Stage{1, {Src{1, 16, 4}, Src{0, 0, 0}}, 0, 0}, Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0},
                             Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0}}, 2, 16, 1};  // 64 -> 32
    default: return Group{1, {Stage{1, {Src{1, 16, 4}, Src{0, 0, 0}}, 0, 1}, Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0},
                              Stage{0, {Src{0, 0, 0}, Src{0, 0, 0}}, 0, 0}}, 2, 16, 1};  // conv7
  }
}

// ---- compile-time geometry of a group ---------------------------------------------------------
__host__ __device__ constexpr int stage_chunks(const Stage& s) { return s.src[0].chunks + (s.nsrc > 1 ? s.src[1].chunks : 0); }
__host__ __device__ constexpr int stage_wbytes(const Group& g, int i) {
  const Stage& s = g.st[i];
  if (g.split) return s.first ? 4 * kBTile : 2 * 3 * (stage_chunks(s) / 4) * kBTile;  // w_hi tiles then w_lo tiles
  return s.first ? 2 * kBTile : 3 * (stage_chunks(s) / 2) * kBTile;
}
__host__ __device__ constexpr int ring_bytes(const Src& r) { return r.chunks * kRowBytes * r.depth; }
__host__ __device__ constexpr int w_offset(const Group& g, int s) {
  int o = 0;
  for (int i = 0; i < s; ++i) o += stage_wbytes(g, i);
  return o;
}
__host__ __device__ constexpr int ring_index(const Group& g, int s, int j) {  // rings enumerated (stage, src)
  int o = 0;
  for (int i = 0; i < s; ++i) o += g.st[i].nsrc;
  return o + j;
}
__host__ __device__ constexpr int n_rings(const Group& g) { return ring_index(g, g.S, 0); }
__host__ __device__ constexpr int ring_offset(const Group& g, int s, int j) {
  int o = w_offset(g, g.S);
  for (int i = 0; i < g.S; ++i)
    for (int k = 0; k < g.st[i].nsrc; ++k) {
      if (i == s && k == j) return o;
      o += ring_bytes(g.st[i].src[k]);
    }
  return o;
}
__host__ __device__ constexpr int bar_offset(const Group& g) { return (ring_offset(g, g.S, 0) + 15) & ~15; }
// barriers: per ring full[depth], empty[depth]; per stage acc_ready[slots], acc_free[slots]; wbar
__host__ __device__ constexpr int ring_bar_index(const Group& g, int s, int j) {
  int o = 0;
  for (int i = 0; i < g.S; ++i)
    for (int k = 0; k < g.st[i].nsrc; ++k) {
      if (i == s && k == j) return o;
      o += 2 * g.st[i].src[k].depth;
    }
  return o;
}
__host__ __device__ constexpr int acc_bar_index(const Group& g, int s) { return ring_bar_index(g, g.S, 0) + 2 * g.slots * s; }
__host__ __device__ constexpr int phys_slots(const Group& g) { return g.slots + (g.ovf ? 2 : 0); }  // per stage, in TMEM
__host__ __device__ constexpr int n_bars(const Group& g) { return acc_bar_index(g, g.S) + 1; }
__host__ __device__ constexpr int smem_bytes(const Group& g) { return bar_offset(g) + 8 * n_bars(g) + 16; }
__host__ __device__ constexpr int producer_index(const Group& g, int s, int j) {  // which producer warp feeds ring (s, j)
  int o = 0;
  for (int i = 0; i < g.S; ++i)
    for (int k = 0; k < g.st[i].nsrc; ++k) {
      if (i == s && k == j) return o;
      o += g.st[i].src[k].ext ? 1 : 0;
    }
  return o;
}
__host__ __device__ constexpr int valid_px(const Group& g) { return kTileX - 2 * (g.S - 1); }
// Epilogue groups (of four warps) per stage; they take the output rows round-robin.  Two.  (With overflow slots a
// slot must always belong to one group: rows go by y % groups and `slots` is a multiple of it.  Three groups were
// measured with the overflow slots -- 896 threads, 72 registers, small spills -- and were slower: conv4-5 +9 %.)
#ifndef LLIE_OVF_GROUPS
#define LLIE_OVF_GROUPS 2
#endif
__host__ __device__ constexpr int epi_groups(const Group& g) { return g.ovf ? LLIE_OVF_GROUPS : 2; }
__host__ __device__ constexpr int n_threads(const Group& g) { return (kFirstEpiWarp + 4 * epi_groups(g) * g.S) * 32; }

struct ChainArgs {
  const __nv_bfloat16* ext[kMaxStages][2][2];  // [stage][src][tensor]: HBM inputs of external rings
  const __nv_bfloat16* wpk[kMaxStages];        // weight tiles of each stage
  float bias[kMaxStages][32];                  // per stage; read as constant-bank operands by the epilogues
  __nv_bfloat16* out[kMaxStages];              // HBM destination of a stage's activation, or nullptr
  const float* x;                              // last stage: the caller's image / output / optional maps
  float* y;
  float* maps;
  ImgView vx, vm;
  int iters;
  float sign;
  int u8;  // sample type of x / y: 0 = float, 1 = uint8_t (value = byte / 255; stored as rint(255 * value)), 2 = uint16_t (65535), 3 = bfloat16
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
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

// Rows a unit touches at stage s (L = last stage index): outputs [ys, ye), inputs [in_lo, in_hi).
struct StageRows { int ys, ye, in_lo, in_hi; };
__device__ __forceinline__ StageRows stage_rows(int ys, int ye, int H, int L, int s) {
  StageRows r;
  r.ys = max(0, ys - (L - s));
  r.ye = min(H, ye + (L - s));
  r.in_lo = max(0, r.ys - 1);
  r.in_hi = min(H, r.ye + 1);
  return r;
}

// ---- shared-memory view: every offset is a template constant ------------------------------------
template <int I> using IC = std::integral_constant<int, I>;

template <int GID>
struct Smem {
  uint8_t* base;
  template <int s> __device__ __forceinline__ uint8_t* weights() const {
    constexpr int o = w_offset(group_cfg(GID), s);
    return base + o;
  }
  template <int s, int j> __device__ __forceinline__ uint8_t* ring() const {
    constexpr int o = ring_offset(group_cfg(GID), s, j);
    return base + o;
  }
  __device__ __forceinline__ uint64_t* bars() const {
    constexpr int o = bar_offset(group_cfg(GID));
    return reinterpret_cast<uint64_t*>(base + o);
  }
  template <int s, int j> __device__ __forceinline__ uint64_t* ring_full() const {
    constexpr int i = ring_bar_index(group_cfg(GID), s, j);
    return bars() + i;
  }
  template <int s, int j> __device__ __forceinline__ uint64_t* ring_empty() const {
    constexpr int i = ring_bar_index(group_cfg(GID), s, j) + group_cfg(GID).st[s].src[j].depth;
    return bars() + i;
  }
  template <int s> __device__ __forceinline__ uint64_t* acc_ready() const {
    constexpr int i = acc_bar_index(group_cfg(GID), s);
    return bars() + i;
  }
  template <int s> __device__ __forceinline__ uint64_t* acc_free() const {
    constexpr int i = acc_bar_index(group_cfg(GID), s) + group_cfg(GID).slots;
    return bars() + i;
  }
  __device__ __forceinline__ uint64_t* wbar() const {
    constexpr int i = n_bars(group_cfg(GID)) - 1;
    return bars() + i;
  }
  __device__ __forceinline__ uint32_t* tmem_ptr() const {
    constexpr int i = n_bars(group_cfg(GID));
    return reinterpret_cast<uint32_t*>(bars() + i);
  }
};

// Calls f(IC<s>, IC<j>) for every ring (stage s, source j) of the group, fully unrolled.
template <int GID, class F>
__device__ __forceinline__ void for_each_ring(F&& f) {
  constexpr Group G = group_cfg(GID);
  if constexpr (G.S > 0 && G.st[0].nsrc > 0) f(IC<0>{}, IC<0>{});
  if constexpr (G.S > 0 && G.st[0].nsrc > 1) f(IC<0>{}, IC<1>{});
  if constexpr (G.S > 1 && G.st[1].nsrc > 0) f(IC<1>{}, IC<0>{});
  if constexpr (G.S > 1 && G.st[1].nsrc > 1) f(IC<1>{}, IC<1>{});
  if constexpr (G.S > 2 && G.st[2].nsrc > 0) f(IC<2>{}, IC<0>{});
  if constexpr (G.S > 2 && G.st[2].nsrc > 1) f(IC<2>{}, IC<1>{});
}
template <int GID, class F>
__device__ __forceinline__ void for_each_stage(F&& f) {
  constexpr Group G = group_cfg(GID);
  f(IC<0>{});
  if constexpr (G.S > 1) f(IC<1>{});
  if constexpr (G.S > 2) f(IC<2>{});
}

struct Unit { int n, sx, ys, ye; };
__device__ __forceinline__ Unit decode_unit(const ChainArgs& a, int u) {
  Unit r;
  const int seg = u % a.segs, t = u / a.segs;
  r.sx = t % a.strips;
  r.n = t / a.strips;
  r.ys = seg * a.seg_rows;
  r.ye = min(r.ys + a.seg_rows, a.H);
  return r;
}

// ---- producer: one warp per external ring -------------------------------------------------------
template <int GID, int S_, int J_>
__device__ __forceinline__ void producer_loop(const ChainArgs& a, const Smem<GID>& sm) {
  constexpr Group G = group_cfg(GID);
  constexpr Src R = G.st[S_].src[J_];
  constexpr int L = G.S - 1;
  constexpr int kRingRow = R.chunks * kRowBytes;
  constexpr int kPlanes = R.chunks == 1 ? 1 : (G.split ? 8 : 4);  // chunk planes per row of one source tensor
  constexpr int kDepth = R.depth;
  constexpr int kValid = valid_px(G);
  uint8_t* ring = sm.template ring<S_, J_>();
  uint64_t* full = sm.template ring_full<S_, J_>();
  uint64_t* empty = sm.template ring_empty<S_, J_>();
  uint32_t cnt = 0;
  const size_t row_stride = (size_t)kPlanes * a.Wp * 8;  // elements between consecutive rows of a source
  for (int u = blockIdx.x; u < a.units; u += gridDim.x) {
    const Unit un = decode_unit(a, u);
    const StageRows sr = stage_rows(un.ys, un.ye, a.H, L, S_);
    // ring position i holds image pixel x = X0 - 1 + i, X0 = sx * valid - (L - s): activation column X0 - 1 + kPadL
    const int col0 = un.sx * kValid - (L - S_) - 1 + kPadL;
    size_t off = ((size_t)un.n * a.H + sr.in_lo) * row_stride + (size_t)col0 * 8;
    for (int y = sr.in_lo; y < sr.in_hi; ++y, off += row_stride, ++cnt) {
      const uint32_t slot = cnt % kDepth;
      mbar_wait(&empty[slot], ((cnt / kDepth) & 1) ^ 1);
      if (elect_one()) {
        mbar_expect_tx(&full[slot], kRingRow);
        uint8_t* dst = ring + slot * kRingRow;
#pragma unroll
        for (int c = 0; c < R.chunks; ++c) {
          const __nv_bfloat16* src = a.ext[S_][J_][c / kPlanes];
          bulk_g2s(dst + c * kRowBytes, src + off + (size_t)(c % kPlanes) * a.Wp * 8, kRowBytes, &full[slot]);
        }
      }
      __syncwarp();
    }
  }
}

// ---- MMA issue of stage S_: a software-pipelined walk over (unit, input row) -----------------------
// oc = output rows of this stage started before the current unit (position in the TMEM slot ring);
// rc = ring rows consumed so far (the same for every source of a stage); r = absolute input row.
struct RowIt { int u; StageRows sr; int r; uint32_t oc, rc; bool valid; uint32_t fmask; };
// (ovf) fmask bit j = parity of the next wait on acc_free[j]; a slot's parity flips each time a row first touches it.

// (ovf) output rows input row r touches for the first time: all of its rows for the first input row of a unit,
// otherwise the row below it.
__device__ __forceinline__ void first_touch(const StageRows& sr, int r, int& a, int& b) {
  a = (r == sr.in_lo) ? max(r - 1, sr.ys) : r + 1;
  b = min(r + 1, sr.ye - 1);
}

template <int GID, int S_>
__device__ __forceinline__ void row_begin_unit(const ChainArgs& a, RowIt& it) {
  constexpr int L = group_cfg(GID).S - 1;
  it.valid = it.u < a.units;
  if (it.valid) {
    const Unit un = decode_unit(a, it.u);
    it.sr = stage_rows(un.ys, un.ye, a.H, L, S_);
    it.r = it.sr.in_lo;
  }
}
template <int GID, int S_>
__device__ __forceinline__ void row_advance(const ChainArgs& a, RowIt& it) {
  if constexpr (group_cfg(GID).ovf != 0) {
    int fa, fb;
    first_touch(it.sr, it.r, fa, fb);
    for (int o = fa; o <= fb; ++o) it.fmask ^= 1u << (o % group_cfg(GID).slots);
  }
  ++it.rc;
  if (++it.r == it.sr.in_hi) {
    it.oc += it.sr.ye - it.sr.ys;
    it.u += gridDim.x;
    row_begin_unit<GID, S_>(a, it);
  }
}

// What input row `it` needs before its MMAs may be issued: the accumulator slots it touches for the first time
// drained and re-zeroed since their previous use, and its ring row(s) landed.  kBlock: spin; else probe once.
template <int GID, int S_, bool kBlock>
__device__ __forceinline__ bool row_ready(const Smem<GID>& sm, const RowIt& it) {
  constexpr Group G = group_cfg(GID);
  constexpr Stage T = G.st[S_];
  constexpr int NS = G.slots;
  constexpr int kD0 = T.src[0].depth, kD1 = T.nsrc > 1 ? T.src[1].depth : 1;
  auto check = [](uint64_t* bar, uint32_t parity) -> bool {
    if constexpr (kBlock) { mbar_wait(bar, parity); return true; }
    else return mbar_try_wait(bar, parity);
  };
  const int rows = it.sr.ye - it.sr.ys, rr = it.r - it.sr.ys;
  const int lo = max(rr - 1, 0), hi = min(rr + 1, rows - 1);
  uint64_t* acc_free = sm.template acc_free<S_>();
  bool ok = true;
  if constexpr (G.ovf != 0) {
    int fa, fb;
    first_touch(it.sr, it.r, fa, fb);
    for (int o = fa; o <= fb; ++o) ok = check(&acc_free[o % NS], (it.fmask >> (o % NS)) & 1) && ok;
  } else {
    for (int o = (it.r == it.sr.in_lo) ? lo : rr + 1; o <= hi; ++o) {
      const uint32_t og = it.oc + o;
      ok = check(&acc_free[og % NS], (og / NS) & 1) && ok;
    }
  }
  ok = check(&sm.template ring_full<S_, 0>()[it.rc % kD0], (it.rc / kD0) & 1) && ok;
  if constexpr (T.nsrc > 1) ok = check(&sm.template ring_full<S_, (T.nsrc > 1 ? 1 : 0)>()[it.rc % kD1], (it.rc / kD1) & 1) && ok;
  return ok;
}

// Issues the MMAs of input row `cur` (whose barriers have been waited on).  Between the first third of the
// MMAs and the rest, the barriers of the NEXT row are probed (and waited on after the last MMA if they were
// not ready): the mbarrier round trips and the address arithmetic run while the tensor pipe drains its
// queue, instead of between rows with the pipe idle (ncu: the issuing warp spent ~40 % of a row on this).
template <int GID, int S_>
__device__ __forceinline__ void issue_row(const Smem<GID>& sm, uint32_t tmem, const RowIt& cur, const RowIt& nxt) {
  constexpr Group G = group_cfg(GID);
  constexpr Stage T = G.st[S_];
  constexpr int NS = G.slots;
  constexpr int kNsrc = T.nsrc;
  constexpr int kD0 = T.src[0].depth, kD1 = kNsrc > 1 ? T.src[1].depth : 1;
  constexpr int kRing0 = T.src[0].chunks * kRowBytes, kRing1 = kNsrc > 1 ? T.src[1].chunks * kRowBytes : 0;
  constexpr bool kFirst = T.first != 0;
  constexpr bool kSplit = G.split != 0;
  constexpr int K16 = stage_chunks(T) / (kSplit ? 4 : 2);  // K = 16 blocks (of real channels) over all sources
  constexpr int k0 = T.src[0].chunks / (kSplit ? 4 : 2);   // ... of which the first source holds k0
  const StageRows& sr = cur.sr;
  const int r = cur.r;
  const uint32_t ocnt = cur.oc, rc = cur.rc;
  const int rows = sr.ye - sr.ys;
  const int rr = r - sr.ys;                                      // relative output row centred on this input row
  const int lo = max(rr - 1, 0), hi = min(rr + 1, rows - 1);     // relative output rows it contributes to
  uint64_t* acc_ready = sm.template acc_ready<S_>();
  const uint32_t slot0 = rc % kD0, slot1 = rc % kD1;
  tc_fence_after();
  if (elect_one()) {
    const uint64_t b_desc0 = smem_desc(smem_u32(sm.template weights<S_>()), kBPlane, 128);
    const uint64_t a0 = smem_desc(smem_u32(sm.template ring<S_, 0>()), kFirst ? 16 : kRowBytes, 128) +
                        (uint64_t)(slot0 * (kRing0 >> 4));
    uint64_t a1 = 0;
    if constexpr (kNsrc > 1)
      a1 = smem_desc(smem_u32(sm.template ring<S_, (kNsrc > 1 ? 1 : 0)>()), kRowBytes, 128) + (uint64_t)(slot1 * (kRing1 >> 4));
    // cnt output rows starting at slot s0; dyslot of the first: 0 = dy+1 (output above), 1 = dy 0, 2 = dy-1.
    // part 0 = the horizontal tap dx = 0, part 1 = dx = 1, 2 (conv1: its first / remaining MMAs).
    auto issue = [&](uint32_t s0, int cnt, int dyslot, int part) {
      const uint32_t idesc = idesc_bf16_m128(32 * cnt);
      const uint32_t d = tmem + (S_ * phys_slots(G) + s0) * 32;
      const uint64_t b_row = b_desc0 + (uint64_t)(dyslot * (512 >> 4));
      if constexpr (kFirst) {
        if (part == 0) {
          mma_bf16_ss(d, a0, b_row, idesc, 1u);                                // dx = 0, 1
        } else {
          mma_bf16_ss(d, a0 + 2, b_row + (uint64_t)(kBTile >> 4), idesc, 1u);  // dx = 2, (zero)
          if constexpr (kSplit) {  // tiles 0/1 hold w_hi against the hi AND lo channels; tiles 2/3 w_lo against hi
            mma_bf16_ss(d, a0, b_row + (uint64_t)((2 * kBTile) >> 4), idesc, 1u);
            mma_bf16_ss(d, a0 + 2, b_row + (uint64_t)((3 * kBTile) >> 4), idesc, 1u);
          }
        }
      } else {
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
          if ((dx == 0) != (part == 0)) continue;
#pragma unroll
          for (int kk = 0; kk < K16; ++kk) {
            const uint64_t ab = kk < k0 ? a0 : a1;
            const int kl = kk < k0 ? kk : kk - k0;
            if constexpr (kSplit) {
              // a source ring row holds, per tensor, chunk planes hi 0..3 then lo 4..7
              const int hi_chunk = (kl >> 1) * 8 + (kl & 1) * 2;
              const uint64_t a_hi = ab + (uint64_t)((hi_chunk * kRowBytes + dx * 16) >> 4);
              const uint64_t a_lo = a_hi + (uint64_t)((4 * kRowBytes) >> 4);
              const uint64_t b_hi = b_row + (uint64_t)(((dx * K16 + kk) * kBTile) >> 4);
              const uint64_t b_lo = b_hi + (uint64_t)((3 * K16 * kBTile) >> 4);
              mma_bf16_ss(d, a_hi, b_hi, idesc, 1u);
              mma_bf16_ss(d, a_lo, b_hi, idesc, 1u);
              mma_bf16_ss(d, a_hi, b_lo, idesc, 1u);
            } else {
              const uint64_t ad = ab + (uint64_t)((kl * 2 * kRowBytes + dx * 16) >> 4);
              const uint64_t bd = b_row + (uint64_t)(((dx * K16 + kk) * kBTile) >> 4);
              mma_bf16_ss(d, ad, bd, idesc, 1u);
            }
          }
        }
      }
    };
    // ovf: physical slot r % NS + dyslot, three consecutive slots never wrap; else the ring may wrap (two MMAs)
    const uint32_t s0 = G.ovf ? (uint32_t)(r % NS) + (uint32_t)(lo - (rr - 1)) : (ocnt + lo) % NS;
    const int cnt = hi - lo + 1, c1 = G.ovf ? cnt : min(cnt, NS - (int)s0), dyslot = lo - (rr - 1);
    issue(s0, c1, dyslot, 0);
    if (c1 < cnt) issue(0, cnt - c1, dyslot + c1, 0);  // the slot ring wrapped
    // short rows (<= 6 MMAs) gain nothing from the early probe (measured: conv1-3 chain 6.8 -> 7.2 ms with it)
    constexpr bool kProbe = !kFirst && 3 * K16 * (kSplit ? 3 : 1) > 6;
    const bool nxt_ready = !nxt.valid ? true : (kProbe ? row_ready<GID, S_, false>(sm, nxt) : false);
    issue(s0, c1, dyslot, 1);
    if (c1 < cnt) issue(0, cnt - c1, dyslot + c1, 1);
    mma_commit(&sm.template ring_empty<S_, 0>()[slot0]);
    if constexpr (kNsrc > 1) mma_commit(&sm.template ring_empty<S_, (kNsrc > 1 ? 1 : 0)>()[slot1]);
    // output row rr-1 has now received its last tap; at the bottom image border so has row rr
    if (rr >= 1) mma_commit(&acc_ready[G.ovf ? (uint32_t)((r - 1) % NS) : (ocnt + rr - 1) % NS]);
    if (r == sr.in_hi - 1 && rr == rows - 1) mma_commit(&acc_ready[G.ovf ? (uint32_t)(r % NS) : (ocnt + rr) % NS]);
    if (!nxt_ready) row_ready<GID, S_, true>(sm, nxt);
  }
  __syncwarp();
}

// ---- epilogue of stage S_: four warps, one TMEM lane quadrant each ---------------------------------
template <int GID, int S_>
__device__ __forceinline__ void epilogue_loop(const ChainArgs& a, const Smem<GID>& sm, uint32_t tmem, int g, int q, int lane) {
  constexpr Group G = group_cfg(GID);
  constexpr int L = G.S - 1, NS = G.slots;
  constexpr bool kLast = G.st[S_].last != 0;
  constexpr bool kFeedsNext = S_ < L;                 // writes the internal ring of stage S_+1 (its last source)
  constexpr int SN = kFeedsNext ? S_ + 1 : S_;        // (clamped so that the constants below stay in range)
  constexpr int JN = kFeedsNext ? G.st[SN].nsrc - 1 : 0;
  constexpr int DN = kFeedsNext ? G.st[SN].src[JN].depth : 1;
  constexpr int kValid = valid_px(G);
  uint64_t* acc_ready = sm.template acc_ready<S_>();
  uint64_t* acc_free = sm.template acc_free<S_>();
  const uint32_t lane_base = (uint32_t)(q * 32) << 16;
  constexpr bool kOvfS = G.ovf != 0;
  const uint32_t tcol = tmem + lane_base + S_ * phys_slots(G) * 32;
  // Accumulator slots start from zero (the bias is added below, from the constant bank); a drained
  // slot is re-zeroed with a single-register tcgen05.st.  Group 0 initialises every slot once.
  if (g == 0) {
    for (int s = 0; s < phys_slots(G); ++s) tmem_fill32(tcol + s * 32, 0u);
    tmem_wait_st();
    tc_fence_before();
    __syncwarp();
    if (lane == 0)
      for (int s = 0; s < NS; ++s) mbar_arrive(&acc_free[s]);
  }
  const int m = q * 32 + lane;  // this thread's pixel of the 128-pixel tile
  uint32_t ocnt = 0;            // running output-row counter (same sequence as the issuer's)
  uint32_t rmask = 0;           // (ovf) bit j = parity of this group's next wait on acc_ready[j] (its own slots only)
  for (int u = blockIdx.x; u < a.units; u += gridDim.x) {
    const Unit un = decode_unit(a, u);
    const StageRows sr = stage_rows(un.ys, un.ye, a.H, L, S_);
    const int x = un.sx * kValid - (L - S_) + m;  // image column of this thread's pixel
    const bool in_img = x >= 0 && x < a.W;
    // the strip owns kValid columns of the last stage's output; earlier stages own the same columns
    const bool owned = in_img && m >= (L - S_) && m < (L - S_) + kValid;
    const int rows = sr.ye - sr.ys;
    // plain ring: group g owns the output rows with (counter & 1) == g; ovf: the rows with y % EG == g
    constexpr int EG = epi_groups(G);
    static_assert(!kOvfS || NS % EG == 0, "a slot must always belong to the same epilogue group");
    for (int i = kOvfS ? ((g - sr.ys % EG + EG) % EG) : ((g ^ ocnt) & 1); i < rows; i += EG) {
      const uint32_t og = ocnt + i;
      const int y = sr.ys + i, s = kOvfS ? y % NS : og % NS;
      uint32_t v[32];
      if constexpr (kOvfS) {
        const uint32_t par = (rmask >> s) & 1;
        rmask ^= 1u << s;
        mbar_wait(&acc_ready[s], par);
        tc_fence_after();
        tmem_ld32(tcol + (s + 1) * 32, v);
        if (s == 0 || s == NS - 1) {
          // the row's first tap (s == 0) / last tap (s == NS-1) went to the overflow slot: add the partial sums
          const uint32_t xcol = tcol + (s == 0 ? NS + 1 : 0) * 32;
          uint32_t w[32];
          tmem_ld32(xcol, w);
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(w[j]));
          tmem_fill32(xcol, 0u);
        } else {
          tmem_wait_ld();
        }
        tmem_fill32(tcol + (s + 1) * 32, 0u);
      } else {
        mbar_wait(&acc_ready[s], (og / NS) & 1);
        tc_fence_after();
        tmem_ld32(tcol + s * 32, v);
        tmem_wait_ld();
        tmem_fill32(tcol + s * 32, 0u);
      }
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_free[s]);
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + a.bias[S_][j]);
      if constexpr (!kLast && G.split != 0) {
        // hi = bf16(relu(v)), lo = bf16(relu(v) - hi): 8 chunk planes per pixel (hi 0..3, lo 4..7)
        uint4 o[8];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t hw[4], lw[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float f0 = fmaxf(__uint_as_float(v[8 * c + 2 * k]), 0.f), f1 = fmaxf(__uint_as_float(v[8 * c + 2 * k + 1]), 0.f);
            const __nv_bfloat162 h2 = __floats2bfloat162_rn(f0, f1);
            const __nv_bfloat162 l2 = __floats2bfloat162_rn(f0 - __low2float(h2), f1 - __high2float(h2));
            hw[k] = *reinterpret_cast<const uint32_t*>(&h2);
            lw[k] = *reinterpret_cast<const uint32_t*>(&l2);
          }
          o[c] = make_uint4(hw[0], hw[1], hw[2], hw[3]);
          o[4 + c] = make_uint4(lw[0], lw[1], lw[2], lw[3]);
        }
        if constexpr (kFeedsNext) {
          const uint32_t rs = og % DN;
          mbar_wait(&sm.template ring_empty<SN, JN>()[rs], ((og / DN) & 1) ^ 1);
          uint8_t* dst = sm.template ring<SN, JN>() + rs * (8 * kRowBytes) + m * 16;
          const uint4 z = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
          for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(dst + c * kRowBytes) = in_img ? o[c] : z;
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.template ring_full<SN, JN>()[rs]);
        }
        if (a.out[S_] != nullptr && owned) {
          uint4* dst = reinterpret_cast<uint4*>(a.out[S_] + ((((size_t)un.n * a.H + y) * 8) * a.Wp + (x + kPadL)) * 8);
#pragma unroll
          for (int c = 0; c < 8; ++c) dst[(size_t)c * a.Wp] = o[c];
        }
      } else if constexpr (!kLast) {
        uint4 o[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          o[c].x = pack_bf16_relu(v[8 * c + 0], v[8 * c + 1]);
          o[c].y = pack_bf16_relu(v[8 * c + 2], v[8 * c + 3]);
          o[c].z = pack_bf16_relu(v[8 * c + 4], v[8 * c + 5]);
          o[c].w = pack_bf16_relu(v[8 * c + 6], v[8 * c + 7]);
        }
        if constexpr (kFeedsNext) {
          // ring row #og of the next stage's internal source; columns outside the image are zero padding
          const uint32_t rs = og % DN;
          mbar_wait(&sm.template ring_empty<SN, JN>()[rs], ((og / DN) & 1) ^ 1);
          uint8_t* dst = sm.template ring<SN, JN>() + rs * (4 * kRowBytes) + m * 16;
          const uint4 z = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
          for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4*>(dst + c * kRowBytes) = in_img ? o[c] : z;
          fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.template ring_full<SN, JN>()[rs]);
        }
        if (a.out[S_] != nullptr && owned) {
          uint4* dst = reinterpret_cast<uint4*>(a.out[S_] + ((((size_t)un.n * a.H + y) * 4) * a.Wp + (x + kPadL)) * 8);
#pragma unroll
          for (int c = 0; c < 4; ++c) dst[(size_t)c * a.Wp] = o[c];
        }
      } else {
        if (owned) {
          const long long bx = un.n * a.vx.sn + (long long)y * a.vx.sy + (long long)x * a.vx.sx;
          float p[3];
          if (a.u8 == 1) {
#pragma unroll
            for (int c = 0; c < 3; ++c) p[c] = (float)__ldg(reinterpret_cast<const uint8_t*>(a.x) + bx + c * a.vx.sc) * (1.f / 255.f);
          } else if (a.u8 == 2) {
#pragma unroll
            for (int c = 0; c < 3; ++c) p[c] = (float)__ldg(reinterpret_cast<const uint16_t*>(a.x) + bx + c * a.vx.sc) * (1.f / 65535.f);
          } else if (a.u8 == 3) {
#pragma unroll
            for (int c = 0; c < 3; ++c) p[c] = __bfloat162float(__ldg(reinterpret_cast<const __nv_bfloat16*>(a.x) + bx + c * a.vx.sc));
          } else {
#pragma unroll
            for (int c = 0; c < 3; ++c) p[c] = __ldg(a.x + bx + c * a.vx.sc);
          }
          float A[24];
#pragma unroll
          for (int j = 0; j < 24; ++j) A[j] = fast_tanh(__uint_as_float(v[j]));
#pragma unroll
          for (int it = 0; it < 8; ++it)
            if (it < a.iters) {
#pragma unroll
              for (int c = 0; c < 3; ++c) p[c] = fmaf(a.sign * A[3 * it + c] * p[c], 1.f - p[c], p[c]);
            }
          if (a.u8 == 1) {
#pragma unroll
            for (int c = 0; c < 3; ++c)
              reinterpret_cast<uint8_t*>(a.y)[bx + c * a.vx.sc] = (uint8_t)__float2uint_rn(255.f * __saturatef(p[c]));
          } else if (a.u8 == 2) {
#pragma unroll
            for (int c = 0; c < 3; ++c)
              reinterpret_cast<uint16_t*>(a.y)[bx + c * a.vx.sc] = (uint16_t)__float2uint_rn(65535.f * __saturatef(p[c]));
          } else if (a.u8 == 3) {
#pragma unroll
            for (int c = 0; c < 3; ++c) reinterpret_cast<__nv_bfloat16*>(a.y)[bx + c * a.vx.sc] = __float2bfloat16_rn(p[c]);
          } else {
#pragma unroll
            for (int c = 0; c < 3; ++c) a.y[bx + c * a.vx.sc] = p[c];
          }
          if (a.maps) {
            const long long bm = un.n * a.vm.sn + (long long)y * a.vm.sy + (long long)x * a.vm.sx;
#pragma unroll
            for (int j = 0; j < 24; ++j) a.maps[bm + j * a.vm.sc] = A[j];
          }
        }
      }
    }
    ocnt += rows;
  }
}

// ---- the kernel -----------------------------------------------------------------------------------
template <int GID>
__global__ void __launch_bounds__(n_threads(group_cfg(GID)), 1) dce_chain_kernel(const __grid_constant__ ChainArgs a) {
  constexpr Group G = group_cfg(GID);
  constexpr int kS = G.S;
  constexpr int kRingLo = w_offset(G, G.S), kRingHi = bar_offset(G);
  static_assert(G.S * phys_slots(G) * 32 <= 512, "TMEM: stages x slots x 32 columns must fit 512");
  static_assert(!G.ovf || (G.slots >= 4 && G.slots <= 30 && G.slots % 2 == 0), "overflow-slot scheme: an even number of 4..30 logical slots");
  static_assert(smem_bytes(G) <= 227 * 1024, "shared memory budget");
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const Smem<GID> sm{smem_raw};
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ring bytes the epilogues never write (columns 128..131 of internal rings) must still be finite
  for (int i = threadIdx.x * 16 + kRingLo; i < kRingHi; i += blockDim.x * 16)
    *reinterpret_cast<uint4*>(smem_raw + i) = make_uint4(0u, 0u, 0u, 0u);
  if (threadIdx.x == 0) {
    for_each_ring<GID>([&](auto S, auto J) {
      constexpr int s = decltype(S)::value, j = decltype(J)::value;
      constexpr int depth = group_cfg(GID).st[s].src[j].depth;
      constexpr int full_count = group_cfg(GID).st[s].src[j].ext ? 1 : 4;  // one expect_tx arrive, or four epilogue warps
      for (int k = 0; k < depth; ++k) {
        mbar_init(&sm.template ring_full<s, j>()[k], full_count);
        mbar_init(&sm.template ring_empty<s, j>()[k], 1);
      }
    });
    for_each_stage<GID>([&](auto S) {
      constexpr int s = decltype(S)::value;
      for (int k = 0; k < group_cfg(GID).slots; ++k) {
        mbar_init(&sm.template acc_ready<s>()[k], 1);
        mbar_init(&sm.template acc_free<s>()[k], 4);
      }
    });
    mbar_init(sm.wbar(), 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(sm.tmem_ptr(), 512);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *sm.tmem_ptr();

  if (warp < kS) {
    // ---------------- MMA issuers: warp s issues stage s; stages are ordered only through the mbarriers -------
    // (ONE issuing thread per stage keeps the fp32 accumulation order, hence the result, deterministic)
    if (warp == 0) {
      if (elect_one()) {
        constexpr uint32_t wbytes = w_offset(group_cfg(GID), group_cfg(GID).S);
        mbar_expect_tx(sm.wbar(), wbytes);
        for_each_stage<GID>([&](auto S) {
          constexpr int s = decltype(S)::value;
          bulk_g2s(sm.template weights<s>(), a.wpk[s], stage_wbytes(group_cfg(GID), s), sm.wbar());
        });
      }
      __syncwarp();
    }
    mbar_wait(sm.wbar(), 0);
    for_each_stage<GID>([&](auto S) {
      constexpr int s = decltype(S)::value;
      if (warp == s) {
        RowIt cur;
        cur.u = blockIdx.x; cur.oc = 0; cur.rc = 0; cur.r = 0; cur.fmask = 0;
        row_begin_unit<GID, s>(a, cur);
        if (cur.valid) row_ready<GID, s, true>(sm, cur);
        while (cur.valid) {
          RowIt nxt = cur;
          row_advance<GID, s>(a, nxt);
          issue_row<GID, s>(sm, tmem, cur, nxt);
          cur = nxt;
        }
      }
    });
  } else if (warp < kFirstEpiWarp) {
    // ---------------- producers: warp S + p feeds the p-th external ring -------------------------------
    const int p = warp - kS;
    for_each_ring<GID>([&](auto S, auto J) {
      constexpr int s = decltype(S)::value, j = decltype(J)::value;
      if constexpr (group_cfg(GID).st[s].src[j].ext != 0) {
        constexpr int mine = producer_index(group_cfg(GID), s, j);
        static_assert(group_cfg(GID).S + mine < kFirstEpiWarp, "issuer + producer warps must fit below the epilogue warps");
        if (p == mine) producer_loop<GID, s, j>(a, sm);
      }
    });
  } else {
    // ---------------- epilogues: one group of four warps per stage ----------------------------------------
    const int grp = (warp - kFirstEpiWarp) >> 2, q = warp & 3;
    for_each_stage<GID>([&](auto S) {
      constexpr int s = decltype(S)::value;
      if (grp / epi_groups(group_cfg(GID)) == s) epilogue_loop<GID, s>(a, sm, tmem, grp % epi_groups(group_cfg(GID)), q, lane);
    });
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace chain
}  // namespace llie
