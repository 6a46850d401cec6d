// Node-level Linears on the tensor cores (included by step.cu).  Node tensors stay fp32 in HBM (they are
// ~100x smaller than the edge tensors); each GEMM converts its fp32 operand to a bf16 scratch copy
// (SiLU / dropout mask applied during the conversion), packs the weight block and runs the same tcgen05
// kernels as the edge path with an fp32 epilogue.  K or N above 256 is split into 256-wide passes.
// Returns AETHER_E_UNSUPPORTED when the shape is not covered; the caller then uses the fp32 kernels.
namespace aether {
namespace {

struct NodeTc {
  bool on = false;
  void* a16 = nullptr;      // [rows, 256] bf16
  void* g16 = nullptr;      // [rows, 256] bf16 (backward only)
  void* wp = nullptr;       // 256*256*2 bytes
  float* partial = nullptr; size_t partial_floats = 0;
};

inline int tc_chunk(int n) { return n > 256 ? 256 : n; }
inline bool tc_dim_ok(int n) {
  const int c = tc_chunk(n);
  return (c == 64 || c == 128 || c == 256) && n % c == 0;
}

// C[M,N] = tf(A[M,K]) W^T + b + res      W: [N, K] block with leading dimension ldw
int linear_tc(const NodeTc& c, const float* A, int64_t lda, int64_t M, int K, const float* W, int64_t ldw,
              const float* b, int N, float* C, int64_t ldc, int tf, const float* mask, const float* res, int64_t ldr,
              cudaStream_t st) {
  if (!c.on || !tc_dim_ok(K) || !tc_dim_ok(N) || M < 1 || (ldc & 3) || (res && (ldr & 3))) return AETHER_E_UNSUPPORTED;
  const int Kc = tc_chunk(K), Nc = tc_chunk(N);
  for (int k0 = 0; k0 < K; k0 += Kc) {
    AE_CHECK(tc_convert_bf16(A + k0, lda, M, Kc, tf, mask ? mask + k0 : nullptr, c.a16, Kc, st));
    for (int n0 = 0; n0 < N; n0 += Nc) {
      AE_CHECK(tc_pack_weight(W + (size_t)n0 * ldw + k0, ldw, Nc, Kc, Nc, Kc, 0, 0, c.wp, st));
      TcGemmArgs g;
      g.M = M; g.wpack = c.wp;
      if (k0 == 0) { g.bias = b ? b + n0 : nullptr; if (res) { g.res = res + n0; g.ldr = ldr; } }
      else { g.res = C + n0; g.ldr = ldc; }
      g.out_f32 = C + n0; g.ldo32 = ldc; g.n_valid = Nc;
      AE_CHECK(tc_gemm(g, c.a16, Kc, Kc, Nc, TC_TF_NONE, st));
    }
  }
  return 0;
}

// C16[M,N] (bf16) = A[M,K] W^T     (node terms P_s | P_r gathered by the edge GEMM's epilogue)
int linear_tc_bf16out(const NodeTc& c, const float* A, int64_t lda, int64_t M, int K, const float* W, int64_t ldw, int N,
                      void* C16, int64_t ldc16, cudaStream_t st) {
  if (!c.on || !tc_dim_ok(K) || K > 256 || !tc_dim_ok(N) || M < 1) return AETHER_E_UNSUPPORTED;
  const int Nc = tc_chunk(N);
  AE_CHECK(tc_convert_bf16(A, lda, M, K, TF_NONE, nullptr, c.a16, K, st));
  for (int n0 = 0; n0 < N; n0 += Nc) {
    AE_CHECK(tc_pack_weight(W + (size_t)n0 * ldw, ldw, Nc, K, Nc, K, 0, 0, c.wp, st));
    TcGemmArgs g;
    g.M = M; g.wpack = c.wp;
    g.out_bf16 = (__nv_bfloat16*)C16 + n0; g.ldo = ldc16;
    AE_CHECK(tc_gemm(g, c.a16, K, K, Nc, TC_TF_NONE, st));
  }
  return 0;
}

// gx[rows, k_in] = ((gy[rows, n_out] W + res) * silu'(dsilu_src)) * mask (+ gx)     W: [n_out, k_in], ld ldw
int dgrad_tc(const NodeTc& c, const float* gy, int64_t ldgy, int64_t rows, int n_out, const float* W, int64_t ldw,
             int k_in, float* gx, int64_t ldgx, const float* dsilu_src, const float* mask, const float* res,
             int accumulate, cudaStream_t st) {
  if (!c.on || !tc_dim_ok(n_out) || !tc_dim_ok(k_in) || rows < 1 || (ldgx & 3)) return AETHER_E_UNSUPPORTED;
  const int Kc = tc_chunk(n_out), Nc = tc_chunk(k_in);
  if (n_out > Kc && (dsilu_src || mask || accumulate)) return AETHER_E_UNSUPPORTED;
  for (int k0 = 0; k0 < n_out; k0 += Kc) {
    AE_CHECK(tc_convert_bf16(gy + k0, ldgy, rows, Kc, TF_NONE, nullptr, c.a16, Kc, st));
    for (int n0 = 0; n0 < k_in; n0 += Nc) {
      AE_CHECK(tc_pack_weight(W + (size_t)k0 * ldw + n0, ldw, Nc, Kc, Nc, Kc, 1, 0, c.wp, st));
      TcGemmArgs g;
      g.M = rows; g.wpack = c.wp;
      if (k0 == 0) { if (res) { g.res = res + n0; g.ldr = ldgx; } }
      else { g.res = gx + n0; g.ldr = ldgx; }
      if (dsilu_src) { g.dsilu32 = dsilu_src + n0; g.ldd32 = ldgx; }
      if (mask) { g.emask = mask + n0; g.ldm = ldgx; }
      g.accumulate = accumulate;
      g.out_f32 = gx + n0; g.ldo32 = ldgx; g.n_valid = Nc;
      AE_CHECK(tc_gemm(g, c.a16, Kc, Kc, Nc, TC_TF_NONE, st));
    }
  }
  return 0;
}

// dW[n_out, k_in] = gy[rows, n_out]^T tf(a[rows, k_in])
int wgrad_tc(const NodeTc& c, const float* gy, int64_t ldgy, int64_t rows, int n_out, const float* a, int64_t lda,
             int k_in, int tf, const float* mask, float* dW, int64_t lddw, cudaStream_t st) {
  if (!c.on || !tc_dim_ok(n_out) || !tc_dim_ok(k_in) || rows < 1) return AETHER_E_UNSUPPORTED;
  const int Mc = tc_chunk(n_out), Nc = tc_chunk(k_in);
  if (!tc_wgrad_supported(Mc, Nc)) return AETHER_E_UNSUPPORTED;
  for (int m0 = 0; m0 < n_out; m0 += Mc) {
    AE_CHECK(tc_convert_bf16(gy + m0, ldgy, rows, Mc, TF_NONE, nullptr, c.g16, Mc, st));
    for (int n0 = 0; n0 < k_in; n0 += Nc) {
      AE_CHECK(tc_convert_bf16(a + n0, lda, rows, Nc, tf, mask ? mask + n0 : nullptr, c.a16, Nc, st));
      AE_CHECK(tc_wgrad(c.g16, Mc, Mc, c.a16, Nc, Nc, TC_TF_NONE, rows, dW + (size_t)m0 * lddw + n0, lddw, Nc, 0,
                        nullptr, c.partial, c.partial_floats, st));
    }
  }
  return 0;
}

}  // namespace
}  // namespace aether
