// K1 (tensor-core filter) -- placeholder until the tcgen05 kernel lands in this file.
#include "common.cuh"

int sepb_netvlad_tensor_topk(sepb_ctx* ctx, const sepb_db*, const double*, int, int, const uint8_t*, const int32_t*,
                             const int32_t*, int32_t*, double*) {
  return sepb_fail(ctx, SEPB_E_STATE, "SEPB_NN_TENSOR: tcgen05 filter not built in this revision", __FILE__, __LINE__);
}
