// Row-tiled MSRCR of one image across several devices (config 5) -- placeholder until the
// halo-exchange path lands.
#include "llie_common.cuh"

extern "C" int llie_msrcr_tiled(llie_ctx* const*, int, const float* const*, float* const*, const int*, const int*, int,
                                int, const llie_msrcr_params*) {
  llie::set_error("llie_msrcr_tiled: not built yet");
  return LLIE_E_STATE;
}
