// Host interface of the tcgen05 kernels (tc_gemm.cu, tc_wgrad.cu).
#pragma once
#include "common.cuh"
#include "tc_common.cuh"

namespace aether {

enum { TC_TF_NONE = 0, TC_TF_SILU = 1 };

struct TcGemmArgs {
  int64_t M = 0;                       // rows of A / out
  const void* wpack = nullptr;         // packed weight image from tc_pack_weight (N*K*2 bytes)
  const float* bias = nullptr;         // [N] or null
  // bf16-output epilogue only: + gat16[idx_a[m]*ldg + n] + gat16[idx_b[m]*ldg + goff_b + n]
  const __nv_bfloat16* gat16 = nullptr; int64_t ldg = 0;
  const int32_t* idx_a = nullptr; const int32_t* idx_b = nullptr; int64_t goff_b = 0;  // goff_b: element offset of the second plane
  // bf16-output epilogue only: * silu'(dsilu[m*ldd + n])
  const __nv_bfloat16* dsilu = nullptr; int64_t ldd = 0;
  // fp32-output epilogue only (node-level Linears), applied in this order:
  const float* res = nullptr; int64_t ldr = 0;            // + res[m*ldr + n]
  const float* dsilu32 = nullptr; int64_t ldd32 = 0;      // * silu'(dsilu32[m*ldd32 + n])
  const float* emask = nullptr; int64_t ldm = 0;          // * emask[m*ldm + n]
  int accumulate = 0;                                     // out_f32 += result
  __nv_bfloat16* out_bf16 = nullptr; int64_t ldo = 0;
  float* out_f32 = nullptr; int64_t ldo32 = 0; int n_valid = 0;   // fp32 output writes columns < n_valid
};

int tc_sm_count();
bool tc_gemm_supported(int K, int N);
size_t tc_weight_pack_bytes(int K, int N);
// B[n][k] = transpose ? W[k*ldw+n] : W[n*ldw + (kmod ? k%kmod : k)], zero outside [n_valid,k_valid]
int tc_pack_weight(const float* W, int64_t ldw, int N, int K, int n_valid, int k_valid, int transpose, int kmod,
                   void* out, cudaStream_t st);
// The same for up to TC_PACK_MAX_JOBS images in ONE launch (every weight image of a step: forward, transposed for the
// backward, node-level), so that the step packs once instead of once per GEMM.
constexpr int TC_PACK_MAX_JOBS = 64;
struct TcPackJob {
  const float* W; void* out; int32_t ldw, N, K, n_valid, k_valid; int16_t transpose, kmod;
};
int tc_pack_weights_batched(const TcPackJob* jobs, int n_jobs, cudaStream_t st);
// out = epi(tf(A) B^T); A is bf16 [M, K] with leading dimension lda (elements)
int tc_gemm(const TcGemmArgs& g, const void* A, int64_t lda, int K, int N, int tf, cudaStream_t st);

// Chained edge MLP on a CTA pair (tc_chain.cu): out2 = silu(out1) W2^T + b2 with out1 = tf1(A) W1^T + b1
// (+ gat16[idx_a[m]] + gat16[idx_b[m] + goff_b]); out1 is written only when non-null (save for backward).
struct TcChainArgs {
  int64_t M = 0;
  const void* wpack1 = nullptr;        // tc_pack_weight image, N = H rows, K = K1
  const void* wpack2 = nullptr;        // tc_pack_weight image, N = H rows, K = H
  const float* bias1 = nullptr; const float* bias2 = nullptr;
  const __nv_bfloat16* gat16 = nullptr; int64_t ldg = 0;
  const int32_t* idx_a = nullptr; const int32_t* idx_b = nullptr; int64_t goff_b = 0;
  int64_t gat_rows = 0;                // rows of the idx_a plane of gat16 (needed with box_gather)
  int box_gather = 0;                  // senders of (nearly) every 128-row tile span < 128 node rows: feed P_s by TMA boxes
  __nv_bfloat16* out1 = nullptr; int64_t ldo1 = 0;
  __nv_bfloat16* out2 = nullptr; int64_t ldo2 = 0;
  // fp32 second output instead of out2 (node update x = y + W4 silu(W3 y + b3) + b4, locs.py:241):
  //   out2_f32[m][n] = acc2 + b2[n] + res[m][n]  (res may alias out2_f32: accumulation over K halves), and
  //   optionally the same value rounded to bf16 in copy16 (operand of the next tensor-core GEMM)
  float* out2_f32 = nullptr; int64_t ldo2f = 0;
  const float* res = nullptr; int64_t ldr = 0;
  __nv_bfloat16* copy16 = nullptr; int64_t ldc16 = 0;
};
bool tc_chain_supported(int K1, int H);
// Chained data-gradient GEMMs of the backward (tc_chain_bwd.cu): out1 = (A W1 (+ gat16[idx_b[m]])) * silu'(z1),
// out2 = (out1 W2) * silu'(z2); wpack1 / wpack2 are transposed images.
bool tc_chain_bwd_supported(int H);
int tc_chain_bwd(const TcChainArgs& g, const void* A, int64_t lda, const void* z1, int64_t ldz1, const void* z2, int64_t ldz2,
                 int H, cudaStream_t st);
int tc_chain(const TcChainArgs& g, const void* A, int64_t lda, int K1, int H, int tf1, cudaStream_t st);

// fp32 [M, cols] (leading dimension lda) -> bf16 [M, cols] (leading dimension ldo), optionally through
// SiLU (tf = TF_SILU) and a dropout mask (TF_SILU_MASK, mask has the layout of A)
int tc_convert_bf16(const float* A, int64_t lda, int64_t M, int cols, int tf, const float* mask, void* out, int64_t ldo,
                    cudaStream_t st);

// dW[MG, NA] = sum_e G[e, :]^T tf(ACT[e, :]) over E rows; both operands bf16 row-major.
// partial: tc_wgrad_partial_floats(MG, NA) floats of scratch.  fold32: dW[m][j] = P[m][j] + P[m][32+j]
// for j < n_valid (layer-1 hi/lo operand), otherwise columns < n_valid are written.
bool tc_wgrad_supported(int MG, int NA);
size_t tc_wgrad_partial_floats(int MG, int NA);
int tc_wgrad(const void* G, int64_t ldg, int MG, const void* ACT, int64_t lda, int NA, int tf, int64_t E, float* dW,
             int64_t lddw, int n_valid, int fold32, float* db /* [MG] column sums of G, or null */, float* partial,
             size_t partial_floats, cudaStream_t st);

}  // namespace aether
