// Test hooks for the tcgen05 building blocks (declared in include/aether_b200.h, "test hooks").
// They run exactly the kernels the bf16 step uses, on caller-provided buffers.
#include "tc_gemm.cuh"

using namespace aether;

extern "C" int aether_dbg_tc_gemm(const void* A_bf16, int64_t M, int32_t K, int64_t lda, const float* W, int64_t ldw,
                                  int32_t N, int32_t n_valid, int32_t k_valid, int32_t transpose, int32_t kmod,
                                  int32_t silu_on_load, const float* bias, const void* gat_bf16, int64_t ldg,
                                  const int32_t* idx_a, const int32_t* idx_b, int32_t goff_b, const void* dsilu_bf16,
                                  void* out_bf16, float* out_f32, void* wpack_scratch, void* stream) {
  if (!A_bf16 || !W || !wpack_scratch || ((out_bf16 != nullptr) == (out_f32 != nullptr))) return AETHER_E_BADARG;
  if (!tc_gemm_supported(K, N)) return AETHER_E_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  AE_CHECK(tc_pack_weight(W, ldw, N, K, n_valid, k_valid, transpose, kmod, wpack_scratch, st));
  TcGemmArgs g;
  g.M = M; g.wpack = wpack_scratch; g.bias = bias;
  g.gat16 = (const __nv_bfloat16*)gat_bf16; g.ldg = ldg; g.idx_a = idx_a; g.idx_b = idx_b; g.goff_b = goff_b;
  g.dsilu = (const __nv_bfloat16*)dsilu_bf16; g.ldd = N;
  g.out_bf16 = (__nv_bfloat16*)out_bf16; g.ldo = N;
  g.out_f32 = out_f32; g.ldo32 = N; g.n_valid = n_valid;
  return tc_gemm(g, A_bf16, lda, K, N, silu_on_load ? TC_TF_SILU : TC_TF_NONE, st);
}

extern "C" size_t aether_dbg_tc_wgrad_partial_floats(int32_t MG, int32_t NA) { return tc_wgrad_partial_floats(MG, NA); }

extern "C" int aether_dbg_tc_wgrad(const void* G_bf16, int32_t MG, const void* ACT_bf16, int32_t NA,
                                   int32_t silu_on_load, int64_t E, float* dW, int32_t n_valid, int32_t fold32,
                                   float* db, float* partial, size_t partial_floats, void* stream) {
  if (!G_bf16 || !ACT_bf16 || !dW) return AETHER_E_BADARG;
  return tc_wgrad(G_bf16, MG, MG, ACT_bf16, NA, NA, silu_on_load ? TC_TF_SILU : TC_TF_NONE, E, dW,
                  n_valid, n_valid, fold32, db, partial, partial_floats, (cudaStream_t)stream);
}
