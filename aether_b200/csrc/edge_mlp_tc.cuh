// bf16 tcgen05 edge-MLP kernels (edge_mlp_tc.cu).
#pragma once
#include "common.cuh"

namespace aether {
// Fused first+second Linear of GNNLayer.message_fn (nn/state2state/locs/locs.py:206-212,235) over
// receiver-sorted edges.  Returns AETHER_E_UNSUPPORTED when the shape is not covered, in which case
// the caller uses the fp32 path.
int edge_mlp_tc_forward(const AetherDims& d, int layer, const float* in, int64_t ld_in, const float* W1,
                        const float* b1, const float* W2, const float* b2, const float* pspr, const int32_t* send,
                        const int32_t* recv, float* pre1, float* pre2, cudaStream_t st);
}  // namespace aether
