// Forward geometry kernels: field-network input, per-node heading frames, per-edge local-frame
// features, globalize + Euler update, and the caller-side step preparation.
// Compiled with -fmad=false (see geom.cuh).
#include <cuda_bf16.h>

#include "geom.cuh"

namespace aether {

namespace {

// fin[n] = [x, vel, emb[(q+1) truncated]]   nn/state2state/aether.py:123-133
template <int D>
__global__ void field_input_kernel(int64_t Nn, const float* __restrict__ x, const float* __restrict__ vel,
                                   const float* __restrict__ charges, const float* __restrict__ emb,
                                   float* __restrict__ fin) {
  constexpr int W = 2 * D + 16;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Nn * W) return;
  const int64_t n = i / W;
  const int c = (int)(i % W);
  float v;
  if (c < D) v = x[n * D + c];
  else if (c < 2 * D) v = vel[n * D + c - D];
  else {
    long long idx = (long long)(charges[n] + 1.0f);   // .long() truncates toward zero
    idx = idx < 0 ? 0 : (idx > 2 ? 2 : idx);
    v = emb[idx * 16 + (c - 2 * D)];
  }
  fin[i] = v;
}

// experiments/lorentz/main.py:243-246: h = |v| ; edge_attr = [q_i q_j, |x_i - x_j|]
template <int D>
__global__ void caller_prep_kernel(int64_t E, int64_t Nn, const float* __restrict__ x, const float* __restrict__ vel,
                                   const float* __restrict__ q, const int64_t* __restrict__ send,
                                   const int64_t* __restrict__ recv, float* __restrict__ h, float* __restrict__ ea) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < E) {
    const int64_t a = send[t], b = recv[t];
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < D; ++k) {
      const float dd = x[a * D + k] - x[b * D + k];
      s = s + dd * dd;
    }
    ea[t * 2 + 0] = q[a] * q[b];
    ea[t * 2 + 1] = sqrtf(s);
  }
  if (h && t < Nn) {
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < D; ++k) s = s + vel[t * D + k] * vel[t * D + k];
    h[t] = sqrtf(s);
  }
}

template <int D>
__global__ void node_frame_kernel(int64_t Nn, const float* __restrict__ vel, const float* __restrict__ f,
                                  float* __restrict__ R, float* __restrict__ x0) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= Nn) return;
  node_frame_body<D>(n, vel, f, R, x0);
}

template <int D>
__global__ void edge_features_kernel(int64_t E, const int32_t* __restrict__ send, const int32_t* __restrict__ recv,
                                     const int32_t* __restrict__ perm, const float* __restrict__ x,
                                     const float* __restrict__ vel, const float* __restrict__ f,
                                     const float* __restrict__ R, const float* __restrict__ x0,
                                     const float* __restrict__ orig, float* __restrict__ feat, int ldf) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  edge_features_body<D>(e, send, recv, perm, x, vel, f, R, x0, orig, feat, ldf);
}

// bf16 path: the features as the A operand of the layer-1 tcgen05 GEMM, K = 64 = [hi(32) | lo(32)] with
// hi = bf16(v), lo = bf16(v - hi): the product against bf16 weights keeps ~16 bits of the feature.
#ifdef __CUDACC__
template <int D>
__global__ void edge_features_hilo_kernel(int64_t E, const int32_t* __restrict__ send, const int32_t* __restrict__ recv,
                                          const int32_t* __restrict__ perm, const float* __restrict__ x,
                                          const float* __restrict__ vel, const float* __restrict__ f,
                                          const float* __restrict__ R, const float* __restrict__ x0,
                                          const float* __restrict__ orig, uint4* __restrict__ feat16) {
  // 128 threads = 128 edges per block.  Each thread computes its edge's 128-byte row; the rows go through a
  // swizzled shared-memory tile so that the global stores are contiguous (512 bytes per warp instruction)
  // instead of 32 scattered 16-byte pieces.
  __shared__ uint4 tile[128 * 8];
  const int t = threadIdx.x;
  const int64_t e0 = (int64_t)blockIdx.x * 128;
  const int64_t e = e0 + t;
  constexpr int FE = 7 * D + D * (D - 1) / 2 + 2;
  constexpr int LD = (FE + 3) / 4 * 4;
  if (e < E) {
    float out[32];
    edge_features_values<D>(e, send, recv, perm, x, vel, f, R, x0, orig, out);
#pragma unroll
    for (int k = LD; k < 32; ++k) out[k] = 0.0f;
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const __nv_bfloat16 h0 = __float2bfloat16_rn(out[2 * k]), h1 = __float2bfloat16_rn(out[2 * k + 1]);
      const __nv_bfloat16 l0 = __float2bfloat16_rn(out[2 * k] - __bfloat162float(h0));
      const __nv_bfloat16 l1 = __float2bfloat16_rn(out[2 * k + 1] - __bfloat162float(h1));
      hi[k] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
      lo[k] = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      tile[t * 8 + (k ^ (t & 7))] = make_uint4(hi[4 * k], hi[4 * k + 1], hi[4 * k + 2], hi[4 * k + 3]);
      tile[t * 8 + ((4 + k) ^ (t & 7))] = make_uint4(lo[4 * k], lo[4 * k + 1], lo[4 * k + 2], lo[4 * k + 3]);
    }
  }
  __syncthreads();
  uint4* dst = feat16 + (size_t)e0 * 8;   // 128 bytes per edge
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int i = it * 128 + t, row = i >> 3, col = i & 7;
    if (e0 + row < E) dst[i] = tile[row * 8 + (col ^ (row & 7))];
  }
}
#endif

template <int D>
__global__ void globalize_kernel(int64_t Nn, const float* __restrict__ x, const float* __restrict__ R,
                                 const float* __restrict__ pred, float* __restrict__ out) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= Nn) return;
  globalize_body<D>(n, x, R, pred, out);
}

inline unsigned nblk_g(int64_t n, int b) { return (unsigned)((n + b - 1) / b); }

}  // namespace

int launch_field_input(int D, int64_t Nn, const float* x, const float* vel, const float* charges, const float* emb,
                       float* fin, cudaStream_t st) {
  if (Nn <= 0) return 0;
  const int64_t tot = Nn * (2 * D + 16);
  if (D == 2) field_input_kernel<2><<<nblk_g(tot, 256), 256, 0, st>>>(Nn, x, vel, charges, emb, fin);
  else field_input_kernel<3><<<nblk_g(tot, 256), 256, 0, st>>>(Nn, x, vel, charges, emb, fin);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_node_frame(int D, int64_t Nn, const float* vel, const float* f, float* R, float* x0, cudaStream_t st) {
  if (Nn <= 0) return 0;
  if (D == 2) node_frame_kernel<2><<<nblk_g(Nn, 128), 128, 0, st>>>(Nn, vel, f, R, x0);
  else node_frame_kernel<3><<<nblk_g(Nn, 128), 128, 0, st>>>(Nn, vel, f, R, x0);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_edge_features(int D, int64_t E, const int32_t* send, const int32_t* recv, const int32_t* perm,
                         const float* x, const float* vel, const float* f, const float* R, const float* x0,
                         const float* orig, float* feat, int ldf, cudaStream_t st) {
  if (E <= 0) return 0;
  if (ldf != edge_feat_ld(D)) return AETHER_E_BADARG;
  if (D == 2) edge_features_kernel<2><<<nblk_g(E, 128), 128, 0, st>>>(E, send, recv, perm, x, vel, f, R, x0, orig, feat, ldf);
  else edge_features_kernel<3><<<nblk_g(E, 128), 128, 0, st>>>(E, send, recv, perm, x, vel, f, R, x0, orig, feat, ldf);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_edge_features_hilo(int D, int64_t E, const int32_t* send, const int32_t* recv, const int32_t* perm,
                              const float* x, const float* vel, const float* f, const float* R, const float* x0,
                              const float* orig, void* feat16, cudaStream_t st) {
  if (E <= 0) return 0;
  if (D == 2) edge_features_hilo_kernel<2><<<nblk_g(E, 128), 128, 0, st>>>(E, send, recv, perm, x, vel, f, R, x0, orig, (uint4*)feat16);
  else edge_features_hilo_kernel<3><<<nblk_g(E, 128), 128, 0, st>>>(E, send, recv, perm, x, vel, f, R, x0, orig, (uint4*)feat16);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_globalize(int D, int64_t Nn, const float* x, const float* R, const float* pred, float* out,
                     cudaStream_t st) {
  if (Nn <= 0) return 0;
  if (D == 2) globalize_kernel<2><<<nblk_g(Nn, 128), 128, 0, st>>>(Nn, x, R, pred, out);
  else globalize_kernel<3><<<nblk_g(Nn, 128), 128, 0, st>>>(Nn, x, R, pred, out);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_caller_prep(int D, const float* x, const float* vel, const float* q, const int64_t* send,
                       const int64_t* recv, int64_t E, int64_t Nn, float* h, float* ea, cudaStream_t st) {
  const int64_t tot = E > Nn ? E : Nn;
  if (tot <= 0) return 0;
  if (D == 2) caller_prep_kernel<2><<<nblk_g(tot, 256), 256, 0, st>>>(E, Nn, x, vel, q, send, recv, h, ea);
  else caller_prep_kernel<3><<<nblk_g(tot, 256), 256, 0, st>>>(E, Nn, x, vel, q, send, recv, h, ea);
  AE_LAUNCH_CHECK();
  return 0;
}

}  // namespace aether
