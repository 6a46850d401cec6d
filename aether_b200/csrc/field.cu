// Neural field queried at the particles (FieldNetwork.forward, nn/state2state/aether.py:108-134) in ONE launch:
//   fin = [x, v, emb[(q + 1).long()]]  ->  z1 = W0 fin + b0  ->  z2 = W2 silu(z1) + b2  ->  f = W4 silu(z2) + b4
// One thread per particle, the three weight matrices (2D+16 -> 32 -> 32 -> D, 7.6 KB) in shared memory and read as
// warp-wide broadcasts; fin and the two pre-activations are written for the backward.  The layer-by-layer form was four
// launches of the generic fp32 GEMM at K <= 32 (38-110 us each, prologue-bound); on a receiver-partitioned scene the field
// is replicated over all particles on every rank, so it sat on the critical path of every rollout step.
#include "common.cuh"

namespace aether {
namespace {

constexpr int FH = 32, EMB = 16;

template <int D>
__global__ void __launch_bounds__(128) field_mlp_fwd_kernel(int64_t Nn, const float* __restrict__ x, const float* __restrict__ vel,
                                                            const float* __restrict__ charges, const float* __restrict__ emb,
                                                            const float* __restrict__ W0, const float* __restrict__ b0,
                                                            const float* __restrict__ W2, const float* __restrict__ b2,
                                                            const float* __restrict__ W4, const float* __restrict__ b4,
                                                            float* __restrict__ fin, float* __restrict__ z1o,
                                                            float* __restrict__ z2o, float* __restrict__ f) {
  constexpr int FI = 2 * D + EMB;
  __shared__ float sW0[FH * FI], sb0[FH], sW2[FH * FH], sb2[FH], sW4[D * FH], sb4[D], semb[3 * EMB];
  for (int i = threadIdx.x; i < FH * FI; i += blockDim.x) sW0[i] = W0[i];
  for (int i = threadIdx.x; i < FH * FH; i += blockDim.x) sW2[i] = W2[i];
  for (int i = threadIdx.x; i < D * FH; i += blockDim.x) sW4[i] = W4[i];
  for (int i = threadIdx.x; i < 3 * EMB; i += blockDim.x) semb[i] = emb[i];
  if (threadIdx.x < FH) { sb0[threadIdx.x] = b0[threadIdx.x]; sb2[threadIdx.x] = b2[threadIdx.x]; }
  if (threadIdx.x < D) sb4[threadIdx.x] = b4[threadIdx.x];
  __syncthreads();
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= Nn) return;
  float in[FI];
#pragma unroll
  for (int c = 0; c < D; ++c) { in[c] = x[n * D + c]; in[D + c] = vel[n * D + c]; }
  long long idx = (long long)(charges[n] + 1.0f);   // .long() truncates toward zero; clamped for memory safety
  idx = idx < 0 ? 0 : (idx > 2 ? 2 : idx);
#pragma unroll
  for (int c = 0; c < EMB; ++c) in[2 * D + c] = semb[idx * EMB + c];
#pragma unroll
  for (int c = 0; c < FI; ++c) fin[n * FI + c] = in[c];
  float a1[FH];
#pragma unroll
  for (int o = 0; o < FH; ++o) {
    float s = sb0[o];
#pragma unroll
    for (int k = 0; k < FI; ++k) s = fmaf(in[k], sW0[o * FI + k], s);
    z1o[n * FH + o] = s;
    a1[o] = silu_f(s);
  }
  float a2[FH];
#pragma unroll
  for (int o = 0; o < FH; ++o) {
    float s = sb2[o];
#pragma unroll
    for (int k = 0; k < FH; ++k) s = fmaf(a1[k], sW2[o * FH + k], s);
    z2o[n * FH + o] = s;
    a2[o] = silu_f(s);
  }
#pragma unroll
  for (int o = 0; o < D; ++o) {
    float s = sb4[o];
#pragma unroll
    for (int k = 0; k < FH; ++k) s = fmaf(a2[k], sW4[o * FH + k], s);
    f[n * D + o] = s;
  }
}

// Backward of the same network in two launches: one block per 128 particles computes the input gradient per thread and
// the block's share of every weight / bias gradient (outer products reduced over the block's particles through shared
// memory, fixed order), a second kernel sums the block partials in block order -- deterministic, no atomics.
// Layout of one partial row: dW0 [32*FI] | db0 [32] | dW2 [32*32] | db2 [32] | dW4 [D*32] | db4 [D].
template <int D>
__host__ __device__ constexpr int field_partial_len() { return FH * (2 * D + EMB) + FH + FH * FH + FH + D * FH + D; }

template <int D>
__global__ void __launch_bounds__(128) field_mlp_bwd_kernel(int64_t Nn, const float* __restrict__ gf, const float* __restrict__ fin,
                                                            const float* __restrict__ z1, const float* __restrict__ z2,
                                                            const float* __restrict__ W0, const float* __restrict__ W2,
                                                            const float* __restrict__ W4, float* __restrict__ gfin,
                                                            float* __restrict__ partial) {
  constexpr int FI = 2 * D + EMB, PL = field_partial_len<D>();
  __shared__ float sW0[FH * FI], sW2[FH * FH], sW4[D * FH];
  __shared__ float U[128][FH + 1], V[128][FH + 1];
  for (int i = threadIdx.x; i < FH * FI; i += blockDim.x) sW0[i] = W0[i];
  for (int i = threadIdx.x; i < FH * FH; i += blockDim.x) sW2[i] = W2[i];
  for (int i = threadIdx.x; i < D * FH; i += blockDim.x) sW4[i] = W4[i];
  __syncthreads();
  const int t = threadIdx.x;
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + t;
  const bool live = n < Nn;
  float* out = partial + (size_t)blockIdx.x * PL;
  float g[D], gz2[FH], gz1[FH];
#pragma unroll
  for (int o = 0; o < D; ++o) g[o] = live ? gf[n * D + o] : 0.0f;
  // ---- layer 3: f = W4 silu(z2) + b4 ----
#pragma unroll
  for (int k = 0; k < FH; ++k) {
    const float z = live ? z2[n * FH + k] : 0.0f;
    float s = 0.0f;
#pragma unroll
    for (int o = 0; o < D; ++o) s = fmaf(g[o], sW4[o * FH + k], s);
    gz2[k] = live ? s * dsilu_f(z) : 0.0f;
    V[t][k] = live ? silu_f(z) : 0.0f;           // a2
  }
#pragma unroll
  for (int o = 0; o < D; ++o) U[t][o] = g[o];
  __syncthreads();
  {
    float* dW4 = out + FH * FI + FH + FH * FH + FH;
    for (int p = t; p < D * FH + D; p += 128) {
      float s = 0.0f;
      if (p < D * FH) { const int o = p / FH, k = p % FH; for (int m = 0; m < 128; ++m) s = fmaf(U[m][o], V[m][k], s); }
      else { const int o = p - D * FH; for (int m = 0; m < 128; ++m) s += U[m][o]; }
      dW4[p] = s;
    }
  }
  __syncthreads();
  // ---- layer 2: z2 = W2 silu(z1) + b2 ----
#pragma unroll
  for (int k = 0; k < FH; ++k) {
    const float z = live ? z1[n * FH + k] : 0.0f;
    float s = 0.0f;
#pragma unroll
    for (int o = 0; o < FH; ++o) s = fmaf(gz2[o], sW2[o * FH + k], s);
    gz1[k] = live ? s * dsilu_f(z) : 0.0f;
    V[t][k] = live ? silu_f(z) : 0.0f;           // a1
    U[t][k] = gz2[k];
  }
  __syncthreads();
  {
    float* dW2 = out + FH * FI + FH;
    const int k = t % FH;
    for (int o = t / FH; o < FH; o += 4) {
      float s = 0.0f;
      for (int m = 0; m < 128; ++m) s = fmaf(U[m][o], V[m][k], s);
      dW2[o * FH + k] = s;
    }
    if (t < FH) { float s = 0.0f; for (int m = 0; m < 128; ++m) s += U[m][t]; dW2[FH * FH + t] = s; }
  }
  __syncthreads();
  // ---- layer 1: z1 = W0 fin + b0 ----
  float in[FI];
#pragma unroll
  for (int c = 0; c < FI; ++c) in[c] = live ? fin[n * FI + c] : 0.0f;
#pragma unroll
  for (int c = 0; c < FI; ++c) {
    float s = 0.0f;
#pragma unroll
    for (int o = 0; o < FH; ++o) s = fmaf(gz1[o], sW0[o * FI + c], s);
    if (live) gfin[n * FI + c] = s;
    V[t][c] = in[c];
  }
#pragma unroll
  for (int k = 0; k < FH; ++k) U[t][k] = gz1[k];
  __syncthreads();
  {
    float* dW0 = out;
    for (int p = t; p < FH * FI + FH; p += 128) {
      float s = 0.0f;
      if (p < FH * FI) { const int o = p / FI, c = p % FI; for (int m = 0; m < 128; ++m) s = fmaf(U[m][o], V[m][c], s); }
      else { const int o = p - FH * FI; for (int m = 0; m < 128; ++m) s += U[m][o]; }
      dW0[p] = s;
    }
  }
}

struct FieldGradPtrs { float* dst[6]; int len[6]; };

__global__ void field_partial_reduce_kernel(const float* __restrict__ partial, int blocks, int PL, FieldGradPtrs o) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= PL) return;
  float s = 0.0f;
  for (int b = 0; b < blocks; ++b) s += partial[(size_t)b * PL + i];
  int off = i;
#pragma unroll
  for (int q = 0; q < 6; ++q) {
    if (off < o.len[q]) { o.dst[q][off] = s; return; }
    off -= o.len[q];
  }
}

}  // namespace

size_t field_mlp_bwd_partial_floats(int D, int64_t Nn) {
  const size_t blocks = (size_t)((Nn + 127) / 128);
  return blocks * (size_t)(D == 2 ? field_partial_len<2>() : field_partial_len<3>()) + 64;
}

// gfin [Nn, 2D+16] and the six parameter gradients (written, not accumulated) from gf [Nn, D] and the saved activations
int launch_field_mlp_bwd(int D, int64_t Nn, const float* gf, const float* fin, const float* z1, const float* z2, const float* W0,
                         const float* W2, const float* W4, float* gfin, float* dW0, float* db0, float* dW2, float* db2,
                         float* dW4, float* db4, float* partial, size_t partial_floats, cudaStream_t st) {
  if (D != 2 && D != 3) return AETHER_E_BADARG;
  const int FI = 2 * D + EMB;
  const int PL = D == 2 ? field_partial_len<2>() : field_partial_len<3>();
  const int blocks = (int)((Nn + 127) / 128);
  if (Nn > 0 && (!partial || partial_floats < (size_t)blocks * PL)) return AETHER_E_WORKSPACE;
  if (Nn > 0) {
    if (D == 2) field_mlp_bwd_kernel<2><<<blocks, 128, 0, st>>>(Nn, gf, fin, z1, z2, W0, W2, W4, gfin, partial);
    else field_mlp_bwd_kernel<3><<<blocks, 128, 0, st>>>(Nn, gf, fin, z1, z2, W0, W2, W4, gfin, partial);
    AE_LAUNCH_CHECK();
  }
  FieldGradPtrs o;
  o.dst[0] = dW0; o.len[0] = FH * FI; o.dst[1] = db0; o.len[1] = FH; o.dst[2] = dW2; o.len[2] = FH * FH;
  o.dst[3] = db2; o.len[3] = FH; o.dst[4] = dW4; o.len[4] = D * FH; o.dst[5] = db4; o.len[5] = D;
  field_partial_reduce_kernel<<<(PL + 127) / 128, 128, 0, st>>>(partial, Nn > 0 ? blocks : 0, PL, o);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_field_mlp_fwd(int D, int64_t Nn, const float* x, const float* vel, const float* charges, const float* emb,
                         const float* W0, const float* b0, const float* W2, const float* b2, const float* W4, const float* b4,
                         float* fin, float* z1, float* z2, float* f, cudaStream_t st) {
  if (Nn <= 0) return 0;
  const unsigned blocks = (unsigned)((Nn + 127) / 128);
  if (D == 2) field_mlp_fwd_kernel<2><<<blocks, 128, 0, st>>>(Nn, x, vel, charges, emb, W0, b0, W2, b2, W4, b4, fin, z1, z2, f);
  else if (D == 3) field_mlp_fwd_kernel<3><<<blocks, 128, 0, st>>>(Nn, x, vel, charges, emb, W0, b0, W2, b2, W4, b4, fin, z1, z2, f);
  else return AETHER_E_BADARG;
  AE_LAUNCH_CHECK();
  return 0;
}

}  // namespace aether
