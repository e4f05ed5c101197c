// tcgen05 3xTF32 emission kernel (placeholder until the tensor-core path lands).
#include "common.cuh"
int kpms_ar_loglik_tc(const float*, const float*, const float*, int, int, int, int, int, float*,
                      cudaStream_t) {
  return KPMS_ERR_UNSUPPORTED;
}
