// DCE-Net weight storage shared by the fp32 (dce_fp32.cu) and tensor-core (dce_tc.cu) paths.
#pragma once
#include "llie_common.cuh"

namespace llie {

// (Cout, Cin) per layer -- SURVEY.md 9.2
static constexpr int kCout[7] = {32, 32, 32, 32, 32, 32, 24};
static constexpr int kCin[7] = {3, 32, 32, 32, 64, 64, 64};
constexpr size_t kNumParams = 79416;

struct DceWeights {
  float* d_f32 = nullptr;          // per layer [cin_pad][9][32] then bias[32]
  size_t w_off[7], b_off[7];
  __nv_bfloat16* d_bf16 = nullptr;  // tensor-core operand layout, filled by dce_tc.cu
  std::vector<float> host_raw;     // the caller's blob (kept for the bf16 packer)
};

int dce_pack_bf16(llie_ctx* ctx, DceWeights* w);  // dce_tc.cu

}  // namespace llie
