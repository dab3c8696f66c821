// C-ABI entry of the row-tiled MSRCR (config 5: one image split into horizontal shards, one per
// device; BASELINE.json#north_star "tiles rows across GPUs with halo rows exchanged by NVLink P2P
// copies, not NCCL").  Validates the partition, enables peer access between the devices involved
// and hands over to msrcr_tiled_run() in msrcr_pyramid.cu, which owns the cascade.  No reference
// file exists to cite (/root/reference/README.md:1-2).
#include "llie_common.cuh"

#include <set>

using namespace llie;

namespace {
std::mutex g_peer_mu;
uint64_t g_peer_enabled[64];  // bit b of entry a: this process has enabled peer access a -> b
}  // namespace

extern "C" int llie_msrcr_tiled(llie_ctx* const* ctxs, int n_dev, const float* const* in_shard, float* const* out_shard,
                                const int* row0, const int* rows, int h_total, int w, const llie_msrcr_params* params) {
  LLIE_REQUIRE(ctxs && in_shard && out_shard && row0 && rows && params, LLIE_E_ARG, "null argument");
  LLIE_REQUIRE(n_dev >= 1 && n_dev <= 64, LLIE_E_ARG, "n_dev %d out of range [1,64]", n_dev);
  llie_msrcr_params full;
  LLIE_TRY(msrcr_params_normalise(params, &full));
  params = &full;
  LLIE_REQUIRE(params->blur == LLIE_BLUR_PYRAMID, LLIE_E_ARG, "row-tiled MSRCR uses the pyramid blur only");
  LLIE_REQUIRE(params->finish >= LLIE_FINISH_NONE && params->finish <= LLIE_FINISH_PERCENTILE, LLIE_E_ARG, "unknown finish mode %d",
               params->finish);
  LLIE_REQUIRE(params->finish != LLIE_FINISH_PERCENTILE || (params->finish_k >= 0.f && params->finish_k < 50.f), LLIE_E_ARG,
               "finish_k %g out of range [0, 50) percent", params->finish_k);
  LLIE_REQUIRE(params->finish != LLIE_FINISH_MEANSTD || (params->finish_k > 0.f && params->finish_k <= 16.f), LLIE_E_ARG,
               "finish_k %g out of range (0, 16]", params->finish_k);
  LLIE_REQUIRE(params->n_scales >= 1 && params->n_scales <= 3, LLIE_E_ARG, "row-tiled MSRCR supports 1..3 scales (got %d)",
               params->n_scales);
  LLIE_REQUIRE(params->truncate > 0.f && params->truncate <= 8.f && params->in_bias >= 1e-30f && params->in_bias <= 1e30f &&
                   params->in_scale >= 0.f && params->in_scale <= 1e30f,
               LLIE_E_ARG, "bad MSRCR parameters");
  LLIE_REQUIRE(h_total > 0 && w > 0 && (long long)h_total * w <= (1ll << 31), LLIE_E_SHAPE, "bad image size %d x %d",
               h_total, w);
  // the shards must tile [0, h_total) in order; interior boundaries on multiples of 4 (the stride of
  // the first cascade level) and every shard tall enough to serve its neighbours' 16-row raw halo
  int next = 0;
  for (int d = 0; d < n_dev; ++d) {
    LLIE_REQUIRE(ctxs[d] && in_shard[d] && out_shard[d], LLIE_E_ARG, "null ctx or shard pointer for device slot %d", d);
    LLIE_REQUIRE(in_shard[d] != out_shard[d], LLIE_E_ARG, "in-place operation is not supported (shard %d)", d);
    LLIE_REQUIRE(row0[d] == next, LLIE_E_SHAPE, "shard %d starts at row %d, expected %d (shards must tile the image in order)",
                 d, row0[d], next);
    LLIE_REQUIRE(rows[d] >= 64 || n_dev == 1, LLIE_E_SHAPE, "shard %d has %d rows; at least 64 are required", d, rows[d]);
    LLIE_REQUIRE(rows[d] > 0, LLIE_E_SHAPE, "shard %d is empty", d);
    LLIE_REQUIRE(d == 0 || row0[d] % 4 == 0, LLIE_E_SHAPE, "shard %d starts at row %d: interior boundaries must be multiples of 4",
                 d, row0[d]);
    next += rows[d];
  }
  LLIE_REQUIRE(next == h_total, LLIE_E_SHAPE, "shards cover %d rows, image has %d", next, h_total);
  // distinct ctx objects; either all on one device ("virtual devices") or all on distinct devices
  std::set<llie_ctx*> uniq(ctxs, ctxs + n_dev);
  LLIE_REQUIRE((int)uniq.size() == n_dev, LLIE_E_ARG, "every shard needs its own ctx");
  std::set<int> devs;
  for (int d = 0; d < n_dev; ++d) devs.insert(ctxs[d]->device);
  LLIE_REQUIRE(devs.size() == 1 || (int)devs.size() == n_dev, LLIE_E_ARG,
               "ctxs must live on %d distinct devices, or all on one device", n_dev);
  if (devs.size() > 1) {
    // Peer access is a property of the process: pairs enabled by an earlier call are remembered, so the steady state
    // costs no CUDA call at all (8 GPUs: 56 ordered pairs x two runtime calls per llie_msrcr_tiled before).
    int prev = 0;
    LLIE_CUDA_OK(cudaGetDevice(&prev));
    const int rc = [&]() -> int {
      std::lock_guard<std::mutex> lk(g_peer_mu);
      for (int a : devs)
        for (int b : devs) {
          if (a == b || (a < 64 && b < 64 && (g_peer_enabled[a] >> b & 1))) continue;
          int can = 0;
          LLIE_CUDA_OK(cudaDeviceCanAccessPeer(&can, a, b));
          LLIE_REQUIRE(can, LLIE_E_NOPEER,
                       "device %d cannot access device %d peer-to-peer (NVLink P2P is required; there is no fallback)", a, b);
          LLIE_CUDA_OK(cudaSetDevice(a));
          const cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
          if (e == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError();
          else LLIE_REQUIRE(e == cudaSuccess, LLIE_E_NOPEER, "cudaDeviceEnablePeerAccess(%d -> %d) failed: %s", a, b,
                            cudaGetErrorString(e));
          if (a < 64 && b < 64) g_peer_enabled[a] |= 1ull << b;
        }
      return LLIE_OK;
    }();
    cudaSetDevice(prev);  // on every path: a failed check must not leave the caller on another device
    LLIE_TRY(rc);
  }
  return msrcr_tiled_run(ctxs, n_dev, in_shard, out_shard, row0, rows, h_total, w, *params);
}
