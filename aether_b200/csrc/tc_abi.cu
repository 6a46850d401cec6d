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

// Chained edge MLP (tc_chain.cu): out2 = silu(out1) W2^T + b2, out1 = tf1(A) W1^T + b1 (+ gathered rows).
// wpack_scratch: (K1 + H) * H * 2 bytes.  out1 may be null (inference: the intermediate is not stored).
extern "C" int aether_dbg_tc_chain_ex(const void* A_bf16, int64_t M, int32_t K1, int64_t lda, const float* W1, int64_t ldw1,
                                      int32_t k1_valid, int32_t kmod1, const float* W2, int32_t H, int32_t silu_on_load,
                                      const float* bias1, const float* bias2, const void* gat_bf16, int64_t ldg,
                                      const int32_t* idx_a, const int32_t* idx_b, int32_t goff_b, void* out1_bf16,
                                      void* out2_bf16, void* wpack_scratch, int64_t gat_rows, int32_t box_gather, void* stream) {
  if (!A_bf16 || !W1 || !W2 || !wpack_scratch || !out2_bf16) return AETHER_E_BADARG;
  if (!tc_chain_supported(K1, H)) return AETHER_E_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  char* wp2 = (char*)wpack_scratch + tc_weight_pack_bytes(K1, H);
  AE_CHECK(tc_pack_weight(W1, ldw1, H, K1, H, k1_valid, 0, kmod1, wpack_scratch, st));
  AE_CHECK(tc_pack_weight(W2, H, H, H, H, H, 0, 0, wp2, st));
  TcChainArgs g;
  g.M = M; g.wpack1 = wpack_scratch; g.wpack2 = wp2; g.bias1 = bias1; g.bias2 = bias2;
  g.gat16 = (const __nv_bfloat16*)gat_bf16; g.ldg = ldg; g.idx_a = idx_a; g.idx_b = idx_b; g.goff_b = goff_b;
  g.gat_rows = gat_rows; g.box_gather = box_gather;
  g.out1 = (__nv_bfloat16*)out1_bf16; g.ldo1 = H; g.out2 = (__nv_bfloat16*)out2_bf16; g.ldo2 = H;
  return tc_chain(g, A_bf16, lda, K1, H, silu_on_load ? TC_TF_SILU : TC_TF_NONE, st);
}

extern "C" int aether_dbg_tc_chain(const void* A_bf16, int64_t M, int32_t K1, int64_t lda, const float* W1, int64_t ldw1,
                                   int32_t k1_valid, int32_t kmod1, const float* W2, int32_t H, int32_t silu_on_load,
                                   const float* bias1, const float* bias2, const void* gat_bf16, int64_t ldg,
                                   const int32_t* idx_a, const int32_t* idx_b, int32_t goff_b, void* out1_bf16,
                                   void* out2_bf16, void* wpack_scratch, void* stream) {
  return aether_dbg_tc_chain_ex(A_bf16, M, K1, lda, W1, ldw1, k1_valid, kmod1, W2, H, silu_on_load, bias1, bias2, gat_bf16, ldg,
                                idx_a, idx_b, goff_b, out1_bf16, out2_bf16, wpack_scratch, 0, 0, stream);
}

// Chained data-gradient GEMMs (tc_chain_bwd.cu): out1 = (A W1 (+ gat[idx_b[m]])) * silu'(z1), out2 = (out1 W2) * silu'(z2);
// W1, W2 are [H, H] fp32 row-major (out = in W: the transposed images are packed here).  wpack_scratch: 2*H*H*2 bytes.
extern "C" int aether_dbg_tc_chain_bwd(const void* A_bf16, int64_t M, int32_t H, const float* W1, const float* W2,
                                       const void* gat_bf16, int64_t ldg, const int32_t* idx_b, const void* z1_bf16,
                                       const void* z2_bf16, void* out1_bf16, void* out2_bf16, void* wpack_scratch,
                                       void* stream) {
  if ((A_bf16 && !W1) || !W2 || !z1_bf16 || !z2_bf16 || !out1_bf16 || !out2_bf16 || !wpack_scratch) return AETHER_E_BADARG;
  if (!tc_chain_bwd_supported(H)) return AETHER_E_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  char* wp2 = (char*)wpack_scratch + tc_weight_pack_bytes(H, H);
  if (A_bf16) AE_CHECK(tc_pack_weight(W1, H, H, H, H, H, 1, 0, wpack_scratch, st));
  AE_CHECK(tc_pack_weight(W2, H, H, H, H, H, 1, 0, wp2, st));
  TcChainArgs g;
  g.M = M; g.wpack1 = wpack_scratch; g.wpack2 = wp2;
  g.gat16 = (const __nv_bfloat16*)gat_bf16; g.ldg = ldg; g.idx_b = idx_b;
  g.out1 = (__nv_bfloat16*)out1_bf16; g.ldo1 = H; g.out2 = (__nv_bfloat16*)out2_bf16; g.ldo2 = H;
  return tc_chain_bwd(g, A_bf16, H, z1_bf16, H, z2_bf16, H, H, st);
}

// Probe of the TMA row gather (cp.async.bulk.tensor.2d.tile::gather4): 128 rows idx[0..127] of a bf16 table [Nn, ld], columns
// [col0, col0 + 64), land in one 16 KB operand slot exactly like a 128 x 64 box load (128-byte swizzle); the raw slot is
// copied to raw_out (16384 bytes) for the test to decode.  box_rows: the row extent the tensor map is encoded with.
namespace {
__global__ void gather4_probe_kernel(const __grid_constant__ CUtensorMap tm, const int32_t* __restrict__ idx, int col0,
                                     uint8_t* __restrict__ raw_out) {
  extern __shared__ __align__(1024) uint8_t smem_probe[];
  __shared__ uint64_t bar;
  const uint32_t b = aether::tc::smem_u32(&bar), d = aether::tc::smem_u32(smem_probe);
  if (threadIdx.x == 0) {
    aether::tc::mbar_init(b, 1);
    aether::tc::fence_mbar_init();
    aether::tc::mbar_arrive_expect_tx(b, 128 * 128);
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int i = threadIdx.x;
    asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                 :: "r"(d + 512u * i), "l"((uint64_t)&tm), "r"(col0), "r"(idx[4 * i]), "r"(idx[4 * i + 1]), "r"(idx[4 * i + 2]),
                    "r"(idx[4 * i + 3]), "r"(b) : "memory");
  }
  aether::tc::mbar_wait(b, 0);
  for (int k = threadIdx.x; k < 16384 / 16; k += blockDim.x)
    reinterpret_cast<uint4*>(raw_out)[k] = reinterpret_cast<const uint4*>(smem_probe)[k];
}
}  // namespace

extern "C" int aether_dbg_gather4_probe(const void* table_bf16, int64_t n_rows, int32_t n_cols, int64_t ld, int32_t box_rows,
                                        const int32_t* idx128, int32_t col0, void* raw_out, void* stream) {
  if (!table_bf16 || !idx128 || !raw_out) return AETHER_E_BADARG;
  CUtensorMap tm;
  AE_CHECK(tc_make_tmap_bf16(&tm, table_bf16, n_rows, n_cols, ld, box_rows));
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaFuncSetAttribute(gather4_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 + 1024);
  if (e != cudaSuccess) return (int)e;
  gather4_probe_kernel<<<1, 128, 16384, st>>>(tm, idx128, col0, (uint8_t*)raw_out);
  AE_LAUNCH_CHECK();
  return 0;
}
