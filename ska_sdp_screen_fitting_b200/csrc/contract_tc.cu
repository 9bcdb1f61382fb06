// Basis contraction, tcgen05 variant (variant 1) -- placeholder until the TMEM kernel lands:
// reports "not supported" so that variant -1 (auto) takes the FP32 SIMT kernel and an explicit
// variant 1 request fails loudly with SSF_ENOTSUP.
#include "common.cuh"

namespace ssf {
bool contract_tc_supported(bool, int, int64_t, int64_t, const void *, const void *) { return false; }
int contract_tc(bool, const float *, const float *, int, int, int64_t, int64_t, float *, cudaStream_t) {
  set_error("tcgen05 contraction variant is not built into this library");
  return SSF_ENOTSUP;
}
}  // namespace ssf
