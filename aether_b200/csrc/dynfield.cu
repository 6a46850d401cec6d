// Latent field network of DynamicFieldAether (nn/state2state/dynamic_field_aether.py:11-47), forward and backward:
//
//   summary z[b]   = sum_n softmax_n(gate(x_n)) * nn(x_n)         graph_pool.py:6-29 (PyG AttentionalAggregation: the gate
//                                                                  is soft-maxed over the nodes of a graph, the values summed)
//   modulators     = FiLM.modulator(z[b]) -> (gamma, beta)         film.py:40-60, two of them
//   field(x_n)     = L3 silu(film2(L2 silu(film1(L1 [x, v, emb(q)])))) film.py:5-37,  film(y) = (1 + gamma) y + beta
//
// Everything is node- or graph-level and 32 wide, so the Linears run on the generic fp32 GEMM family (gemm_f32.cu: forward
// with SiLU-on-load, data gradient with the silu' epilogue, deterministic split-K weight gradient); the two pieces the
// reference takes from PyG / builds from broadcasting -- the segmented soft-max attention pooling and the per-graph
// feature-wise modulation -- are the kernels below.  Graphs are equal-sized blocks of `num_nodes` consecutive nodes, as
// the reference's reshape(-1, num_nodes, .) assumes (dynamic_field_aether.py:32).  No atomics: every reduction is one
// warp walking its graph in node order.
#include "common.cuh"

namespace aether {
namespace {

constexpr int FH = 32;   // field_hidden_size = summary_dim (dynamic_field_aether.py:65-66)

// parameter slots, state_dict order of `field_net`
enum {
  Q_G0_W = 0, Q_G0_B, Q_G2_W, Q_G2_B, Q_N0_W, Q_N0_B, Q_N2_W, Q_N2_B,
  Q_L1_W = 8, Q_L1_B, Q_L2_W, Q_L2_B, Q_L3_W, Q_L3_B,
  Q_F1 = 14,      // film_1.modulator.{0,2,4}.{weight,bias}
  Q_F2 = 20,      // film_2.modulator.{0,2,4}.{weight,bias}
  Q_EMB = 26, Q_NUM = 27
};

// one warp per graph, lane = summary column
__global__ void attn_pool_fwd_kernel(int64_t B, int n, const float* __restrict__ gate, const float* __restrict__ nv,
                                     float* __restrict__ alpha, float* __restrict__ z) {
  const int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (b >= B) return;
  const float* g = gate + b * n;
  float mx = -INFINITY;
  for (int i = lane; i < n; i += 32) mx = fmaxf(mx, g[i]);
  for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float den = 0.0f;
  for (int i = lane; i < n; i += 32) den += expf(g[i] - mx);
  for (int o = 16; o; o >>= 1) den += __shfl_xor_sync(0xffffffffu, den, o);
  const float inv = 1.0f / (den + 1e-16f);          // PyG softmax: out / (sum + 1e-16)
  float acc = 0.0f;
  for (int i = 0; i < n; ++i) {
    const float a = expf(g[i] - mx) * inv;
    if (lane == 0) alpha[b * n + i] = a;
    acc += a * nv[(b * n + i) * FH + lane];
  }
  z[b * FH + lane] = acc;
}

// g_nv[i,:] = alpha_i gz[b,:] ; g_gate[i] = alpha_i (d_i - sum_m alpha_m d_m), d_i = gz[b] . nv[i]
__global__ void attn_pool_bwd_kernel(int64_t B, int n, const float* __restrict__ alpha, const float* __restrict__ nv,
                                     const float* __restrict__ gz, float* __restrict__ g_nv, float* __restrict__ g_gate) {
  const int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (b >= B) return;
  const float gzc = gz[b * FH + lane];
  float s = 0.0f;
  for (int i = 0; i < n; ++i) {
    float d = gzc * nv[(b * n + i) * FH + lane];
    for (int o = 16; o; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    s += alpha[b * n + i] * d;
  }
  for (int i = 0; i < n; ++i) {
    const float a = alpha[b * n + i];
    float d = gzc * nv[(b * n + i) * FH + lane];
    for (int o = 16; o; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    g_nv[(b * n + i) * FH + lane] = a * gzc;
    if (lane == 0) g_gate[b * n + i] = a * (d - s);
  }
}

// u[i,c] = (1 + gamma[b,c]) y[i,c] + beta[b,c],  mod[b] = [gamma(32) | beta(32)]
__global__ void film_fwd_kernel(int64_t Nn, int n, const float* __restrict__ y, const float* __restrict__ mod,
                                float* __restrict__ u) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Nn * FH) return;
  const int64_t node = i / FH;
  const int c = (int)(i % FH);
  const int64_t b = node / n;
  u[i] = (1.0f + mod[b * 2 * FH + c]) * y[i] + mod[b * 2 * FH + FH + c];
}

// gy = (1 + gamma) gu ; gmod[b] = [sum_i gu y | sum_i gu]   (one warp per graph, lane = column, node order)
__global__ void film_bwd_kernel(int64_t B, int n, const float* __restrict__ gu, const float* __restrict__ y,
                                const float* __restrict__ mod, float* __restrict__ gy, float* __restrict__ gmod) {
  const int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (b >= B) return;
  const float g1 = 1.0f + mod[b * 2 * FH + lane];
  float sg = 0.0f, sb = 0.0f;
  for (int i = 0; i < n; ++i) {
    const int64_t k = (b * n + i) * FH + lane;
    const float g = gu[k];
    sg += g * y[k];
    sb += g;
    gy[k] = g1 * g;
  }
  gmod[b * 2 * FH + lane] = sg;
  gmod[b * 2 * FH + FH + lane] = sb;
}

struct Ws {     // forward activations kept for the backward
  float *fin, *gpre, *gate, *npre, *nv, *alpha, *z, *m1[2], *m2[2], *mod[2], *y1, *u1, *y2, *u2;
  size_t bytes;
};
struct Carve {
  char* base; size_t off = 0;
  explicit Carve(void* p) : base((char*)p) {}
  float* f(size_t n) { float* r = (float*)(base + off); off += align_up(n * sizeof(float), 256); return r; }
};
Ws carve_ws(int D, int64_t Nn, int64_t B, void* base) {
  Ws w;
  Carve c(base);
  w.fin = c.f((size_t)Nn * (2 * D + 16));
  w.gpre = c.f((size_t)Nn * FH); w.gate = c.f((size_t)Nn); w.npre = c.f((size_t)Nn * FH); w.nv = c.f((size_t)Nn * FH);
  w.alpha = c.f((size_t)Nn); w.z = c.f((size_t)B * FH);
  for (int k = 0; k < 2; ++k) { w.m1[k] = c.f((size_t)B * FH); w.m2[k] = c.f((size_t)B * FH); w.mod[k] = c.f((size_t)B * 2 * FH); }
  w.y1 = c.f((size_t)Nn * FH); w.u1 = c.f((size_t)Nn * FH); w.y2 = c.f((size_t)Nn * FH); w.u2 = c.f((size_t)Nn * FH);
  w.bytes = c.off;
  return w;
}
struct Sc {     // backward scratch
  float *gu2, *gy2, *gu1, *gy1, *gfin, *gmod[2], *gm2, *gm1, *gz, *gnv, *ggate, *gnpre, *ggpre, *partial;
  size_t partial_floats, bytes;
};
Sc carve_sc(int D, int64_t Nn, int64_t B, void* base) {
  Sc s;
  Carve c(base);
  const int FI = 2 * D + 16;
  s.gu2 = c.f((size_t)Nn * FH); s.gy2 = c.f((size_t)Nn * FH); s.gu1 = c.f((size_t)Nn * FH); s.gy1 = c.f((size_t)Nn * FH);
  s.gfin = c.f((size_t)Nn * FI);
  for (int k = 0; k < 2; ++k) s.gmod[k] = c.f((size_t)B * 2 * FH);
  s.gm2 = c.f((size_t)B * FH); s.gm1 = c.f((size_t)B * FH); s.gz = c.f((size_t)B * FH);
  s.gnv = c.f((size_t)Nn * FH); s.ggate = c.f((size_t)Nn); s.gnpre = c.f((size_t)Nn * FH); s.ggpre = c.f((size_t)Nn * FH);
  size_t m = 256 * 48 + 64;
  auto up = [&](size_t v) { if (v > m) m = v; };
  up(gemm_tn_partial_floats(2 * FH, FI, Nn)); up(gemm_tn_partial_floats(2 * FH, FH, Nn));
  up(gemm_tn_partial_floats(2 * FH, FH, B)); up(colsum_partial_floats(Nn, 2 * FH)); up(colsum_partial_floats(B, 2 * FH));
  s.partial_floats = m + 64;
  s.partial = c.f(s.partial_floats);
  s.bytes = c.off;
  return s;
}

// C[M,N] = tf(A[M,K]) W^T + b
int lin(const float* A, int64_t lda, int64_t M, int K, const float* W, const float* b, int N, float* C, int64_t ldc, int tf,
        cudaStream_t st) {
  GemmArgs g;
  g.A = A; g.lda = lda; g.B = W; g.ldb = K; g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K; g.tf = tf; g.bias = b;
  return gemm_f32(GEMM_NT, g, st);
}
// backward of lin: dW = gC^T tf(A), db = colsum(gC), gA = (gC W) * silu'(A) when tf (A is then a pre-activation)
int lin_bwd(const float* gC, int64_t ldgc, int64_t M, int N, const float* A, int64_t lda, int K, const float* W, int tf,
            float* dW, float* db, float* gA, int64_t ldga, int accumulate, const Sc& s, cudaStream_t st) {
  GemmArgs t;
  t.A = gC; t.lda = ldgc; t.B = A; t.ldb = lda; t.C = dW; t.ldc = K; t.M = N; t.N = K; t.K = M; t.tf = tf;
  t.partial = s.partial; t.partial_floats = s.partial_floats;
  AE_CHECK(gemm_f32(GEMM_TN, t, st));
  if (db) AE_CHECK(colsum_f32(gC, ldgc, M, N, db, s.partial, s.partial_floats, st));
  if (gA) {
    GemmArgs d;
    d.A = gC; d.lda = ldgc; d.B = W; d.ldb = K; d.C = gA; d.ldc = ldga; d.M = M; d.N = K; d.K = N;
    if (tf == TF_SILU) { d.dsilu = A; d.ldd = lda; }
    d.accumulate = accumulate;
    AE_CHECK(gemm_f32(GEMM_NN, d, st));
  }
  return 0;
}

bool args_ok(int D, int64_t Nn, int num_nodes) {
  return (D == 2 || D == 3) && Nn >= 0 && num_nodes > 0 && Nn % num_nodes == 0 && Nn < ((int64_t)1 << 31);
}

}  // namespace
}  // namespace aether

using namespace aether;

extern "C" size_t aether_dynfield_workspace_bytes(int32_t D, int64_t Nn, int32_t num_nodes) {
  if (!args_ok(D, Nn, num_nodes)) return 0;
  return carve_ws(D, Nn, Nn / num_nodes, nullptr).bytes + 256;
}
extern "C" size_t aether_dynfield_scratch_bytes(int32_t D, int64_t Nn, int32_t num_nodes) {
  if (!args_ok(D, Nn, num_nodes)) return 0;
  return carve_sc(D, Nn, Nn / num_nodes, nullptr).bytes + 256;
}

extern "C" int aether_dynfield_forward(int32_t D, int64_t Nn, int32_t num_nodes, const float* const* P, const float* x,
                                       const float* vel, const float* charges, float* field, void* workspace,
                                       size_t workspace_bytes, void* stream) {
  if (!args_ok(D, Nn, num_nodes) || !P || !field) return AETHER_E_BADARG;
  const int64_t B = Nn / num_nodes;
  if (!workspace || ((uintptr_t)workspace & 255) || workspace_bytes < carve_ws(D, Nn, B, nullptr).bytes) return AETHER_E_WORKSPACE;
  if (Nn == 0) return 0;
  const Ws w = carve_ws(D, Nn, B, workspace);
  cudaStream_t st = (cudaStream_t)stream;
  const int FI = 2 * D + 16, XI = 2 * D;
  AE_CHECK(launch_field_input(D, Nn, x, vel, charges, P[Q_EMB], w.fin, st));     // [x | v | emb(q)]; [x | v] = first 2D columns
  // graph summary
  AE_CHECK(lin(w.fin, FI, Nn, XI, P[Q_G0_W], P[Q_G0_B], FH, w.gpre, FH, TF_NONE, st));
  AE_CHECK(lin(w.gpre, FH, Nn, FH, P[Q_G2_W], P[Q_G2_B], 1, w.gate, 1, TF_SILU, st));
  AE_CHECK(lin(w.fin, FI, Nn, XI, P[Q_N0_W], P[Q_N0_B], FH, w.npre, FH, TF_NONE, st));
  AE_CHECK(lin(w.npre, FH, Nn, FH, P[Q_N2_W], P[Q_N2_B], FH, w.nv, FH, TF_SILU, st));
  attn_pool_fwd_kernel<<<(unsigned)((B * 32 + 255) / 256), 256, 0, st>>>(B, num_nodes, w.gate, w.nv, w.alpha, w.z);
  AE_LAUNCH_CHECK();
  // the two FiLM modulators (per graph)
  for (int k = 0; k < 2; ++k) {
    const int q = k == 0 ? Q_F1 : Q_F2;
    AE_CHECK(lin(w.z, FH, B, FH, P[q], P[q + 1], FH, w.m1[k], FH, TF_NONE, st));
    AE_CHECK(lin(w.m1[k], FH, B, FH, P[q + 2], P[q + 3], FH, w.m2[k], FH, TF_SILU, st));
    AE_CHECK(lin(w.m2[k], FH, B, FH, P[q + 4], P[q + 5], 2 * FH, w.mod[k], 2 * FH, TF_SILU, st));
  }
  // modulated field MLP
  const unsigned eb = (unsigned)((Nn * FH + 255) / 256);
  AE_CHECK(lin(w.fin, FI, Nn, FI, P[Q_L1_W], P[Q_L1_B], FH, w.y1, FH, TF_NONE, st));
  film_fwd_kernel<<<eb, 256, 0, st>>>(Nn, num_nodes, w.y1, w.mod[0], w.u1);
  AE_LAUNCH_CHECK();
  AE_CHECK(lin(w.u1, FH, Nn, FH, P[Q_L2_W], P[Q_L2_B], FH, w.y2, FH, TF_SILU, st));
  film_fwd_kernel<<<eb, 256, 0, st>>>(Nn, num_nodes, w.y2, w.mod[1], w.u2);
  AE_LAUNCH_CHECK();
  AE_CHECK(lin(w.u2, FH, Nn, FH, P[Q_L3_W], P[Q_L3_B], D, field, D, TF_SILU, st));
  return 0;
}

extern "C" int aether_dynfield_backward(int32_t D, int64_t Nn, int32_t num_nodes, const float* const* P,
                                        const float* charges, const float* grad_field, float* const* G, float* grad_x,
                                        float* grad_vel, const void* workspace, size_t workspace_bytes, void* scratch,
                                        size_t scratch_bytes, void* stream) {
  if (!args_ok(D, Nn, num_nodes) || !P || !G || !grad_field) return AETHER_E_BADARG;
  const int64_t B = Nn / num_nodes;
  if (!workspace || ((uintptr_t)workspace & 255) || workspace_bytes < carve_ws(D, Nn, B, nullptr).bytes) return AETHER_E_WORKSPACE;
  if (!scratch || ((uintptr_t)scratch & 255) || scratch_bytes < carve_sc(D, Nn, B, nullptr).bytes) return AETHER_E_WORKSPACE;
  if (Nn == 0) return 0;
  const Ws w = carve_ws(D, Nn, B, const_cast<void*>(workspace));
  const Sc s = carve_sc(D, Nn, B, scratch);
  cudaStream_t st = (cudaStream_t)stream;
  const int FI = 2 * D + 16, XI = 2 * D;
  const unsigned gb = (unsigned)((B * 32 + 255) / 256);
  // field MLP, top down
  AE_CHECK(lin_bwd(grad_field, D, Nn, D, w.u2, FH, FH, P[Q_L3_W], TF_SILU, G[Q_L3_W], G[Q_L3_B], s.gu2, FH, 0, s, st));
  film_bwd_kernel<<<gb, 256, 0, st>>>(B, num_nodes, s.gu2, w.y2, w.mod[1], s.gy2, s.gmod[1]);
  AE_LAUNCH_CHECK();
  AE_CHECK(lin_bwd(s.gy2, FH, Nn, FH, w.u1, FH, FH, P[Q_L2_W], TF_SILU, G[Q_L2_W], G[Q_L2_B], s.gu1, FH, 0, s, st));
  film_bwd_kernel<<<gb, 256, 0, st>>>(B, num_nodes, s.gu1, w.y1, w.mod[0], s.gy1, s.gmod[0]);
  AE_LAUNCH_CHECK();
  AE_CHECK(lin_bwd(s.gy1, FH, Nn, FH, w.fin, FI, FI, P[Q_L1_W], TF_NONE, G[Q_L1_W], G[Q_L1_B], s.gfin, FI, 0, s, st));
  // modulators -> gradient of the graph summary
  for (int k = 0; k < 2; ++k) {
    const int q = k == 0 ? Q_F1 : Q_F2;
    AE_CHECK(lin_bwd(s.gmod[k], 2 * FH, B, 2 * FH, w.m2[k], FH, FH, P[q + 4], TF_SILU, G[q + 4], G[q + 5], s.gm2, FH, 0, s, st));
    AE_CHECK(lin_bwd(s.gm2, FH, B, FH, w.m1[k], FH, FH, P[q + 2], TF_SILU, G[q + 2], G[q + 3], s.gm1, FH, 0, s, st));
    AE_CHECK(lin_bwd(s.gm1, FH, B, FH, w.z, FH, FH, P[q], TF_NONE, G[q], G[q + 1], s.gz, FH, k, s, st));
  }
  // attention pooling
  attn_pool_bwd_kernel<<<gb, 256, 0, st>>>(B, num_nodes, w.alpha, w.nv, s.gz, s.gnv, s.ggate);
  AE_LAUNCH_CHECK();
  AE_CHECK(lin_bwd(s.gnv, FH, Nn, FH, w.npre, FH, FH, P[Q_N2_W], TF_SILU, G[Q_N2_W], G[Q_N2_B], s.gnpre, FH, 0, s, st));
  AE_CHECK(lin_bwd(s.gnpre, FH, Nn, FH, w.fin, FI, XI, P[Q_N0_W], TF_NONE, G[Q_N0_W], G[Q_N0_B], s.gfin, FI, 1, s, st));
  AE_CHECK(lin_bwd(s.ggate, 1, Nn, 1, w.gpre, FH, FH, P[Q_G2_W], TF_SILU, G[Q_G2_W], G[Q_G2_B], s.ggpre, FH, 0, s, st));
  AE_CHECK(lin_bwd(s.ggpre, FH, Nn, FH, w.fin, FI, XI, P[Q_G0_W], TF_NONE, G[Q_G0_W], G[Q_G0_B], s.gfin, FI, 1, s, st));
  // inputs: class embedding and (accumulated) positions / velocities
  AE_CHECK(launch_embedding_bwd(D, Nn, charges, s.gfin, G[Q_EMB], s.partial, s.partial_floats, st));
  if (grad_x || grad_vel) AE_CHECK(launch_add_field_input_grads(D, Nn, s.gfin, grad_x, grad_vel, st));
  return 0;
}
