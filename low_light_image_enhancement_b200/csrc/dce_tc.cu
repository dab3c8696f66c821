// Zero-DCE, bf16 tensor-core path (config 2) -- placeholder until the tcgen05 kernels land.
#include "dce_weights.cuh"

namespace llie {

int dce_pack_bf16(llie_ctx*, DceWeights*) { return LLIE_OK; }

int dce_forward_bf16(llie_ctx*, const float*, float*, float*, int, int, int, int, const llie_dce_params&, cudaStream_t) {
  set_error("bf16 tensor-core path not built yet");
  return LLIE_E_STATE;
}

}  // namespace llie
