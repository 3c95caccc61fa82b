// Placeholder until the tcgen05 kernels land: the tensor-core precisions fail loudly.
#include "common.cuh"
namespace eda {
int mha_tc(const float*, int, float*, int, int, int, int, int, cudaStream_t) {
  set_error("eda_mha_fwd: tcgen05 path not built");
  return EDA_EUNSUPPORTED;
}
}  // namespace eda
