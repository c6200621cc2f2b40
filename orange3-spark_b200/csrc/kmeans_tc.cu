// placeholder until the tcgen05 kernel lands
#include "common.cuh"
bool km_tc_supported(int, int, int64_t) { return false; }
int km_tc_step(const float*, int64_t, int, int64_t, const void*, int, void*, int32_t*, int, bool, cudaStream_t) {
  b2f_set_error("tensor-core KMeans path not built");
  return B2F_EUNSUPPORTED;
}
