#include "edge_mlp_tc.cuh"

namespace aether {
int edge_mlp_tc_forward(const AetherDims&, int, const float*, int64_t, const float*, const float*, const float*,
                        const float*, const float*, const int32_t*, const int32_t*, float*, float*, cudaStream_t) {
  return AETHER_E_UNSUPPORTED;
}
}  // namespace aether
