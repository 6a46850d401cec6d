// HBM-bound edge kernels of the bf16 path: the per-edge tensors [E, H] are stored as bf16
// pre-activations (2H bytes per edge per tensor), reductions accumulate in fp32.
//   segmean_silu_bf16 : y[n] = base[n] + mean_{k in seg(n)} silu(pre2[k])   (locs.py:236-240)
//   segsum_bf16       : out[n] = sum_{k in seg(n)} g[list ? list[k] : k]    (backward of the two gathers)
//   edge_dpre2_bf16   : g2[k] = (ge[k] + gy[recv[k]] / deg) * silu'(pre2[k]) (backward of the mean + SiLU)
//   colsum_bf16       : bias gradients
// Receiver-sorted edges make every segment a contiguous row range: no atomics, fixed summation order.
#include "tc_gemm.cuh"

namespace aether {

namespace {

using tc::bf16_hi;
using tc::bf16_lo;

__device__ __forceinline__ void unpack8(const uint4& u, float (&v)[8]) {
  v[0] = bf16_lo(u.x); v[1] = bf16_hi(u.x); v[2] = bf16_lo(u.y); v[3] = bf16_hi(u.y);
  v[4] = bf16_lo(u.z); v[5] = bf16_hi(u.z); v[6] = bf16_lo(u.w); v[7] = bf16_hi(u.w);
}

// One warp per node.  A row of W bf16 is covered by LPR = W/8 lanes (16 bytes each); the 32/LPR lane
// groups take alternate rows of the segment and are combined with shuffles at the end.
template <int LPR, bool SILU, bool MEAN>
__global__ void segreduce_bf16_kernel(int64_t Nn, const int32_t* __restrict__ ptr, const int32_t* __restrict__ list,
                                      const __nv_bfloat16* __restrict__ src, int64_t lds, const float* base,
                                      int64_t ldb, float* y, int64_t ldy, __nv_bfloat16* y16, int64_t ldy16) {
  constexpr int G = 32 / LPR;
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (n >= Nn) return;
  const int b = ptr[n], e = ptr[n + 1];
  const int grp = lane / LPR, c = (lane % LPR) * 8;
  float s[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = 0.0f;
  int k = b + grp;
  // two rows in flight per lane group
  for (; k + G < e; k += 2 * G) {
    const int64_t r0 = list ? list[k] : k, r1 = list ? list[k + G] : k + G;
    const uint4 u0 = __ldg(reinterpret_cast<const uint4*>(src + (size_t)r0 * lds + c));
    const uint4 u1 = __ldg(reinterpret_cast<const uint4*>(src + (size_t)r1 * lds + c));
    float v0[8], v1[8];
    unpack8(u0, v0);
    unpack8(u1, v1);
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] += SILU ? tc::silu_fast(v0[i]) + tc::silu_fast(v1[i]) : v0[i] + v1[i];
  }
  if (k < e) {
    const int64_t r0 = list ? list[k] : k;
    const uint4 u0 = __ldg(reinterpret_cast<const uint4*>(src + (size_t)r0 * lds + c));
    float v0[8];
    unpack8(u0, v0);
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] += SILU ? tc::silu_fast(v0[i]) : v0[i];
  }
#pragma unroll
  for (int off = LPR; off < 32; off <<= 1)
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] += __shfl_xor_sync(0xffffffffu, s[i], off);
  if (grp == 0) {
    const int deg = e - b;
    const float inv = MEAN ? 1.0f / (float)(deg < 1 ? 1 : deg) : 1.0f;
    float o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = s[i] * inv;
    if (base) {
      const float4 b0 = *reinterpret_cast<const float4*>(base + (size_t)n * ldb + c);
      const float4 b1 = *reinterpret_cast<const float4*>(base + (size_t)n * ldb + c + 4);
      o[0] += b0.x; o[1] += b0.y; o[2] += b0.z; o[3] += b0.w; o[4] += b1.x; o[5] += b1.y; o[6] += b1.z; o[7] += b1.w;
    }
    float4* po = reinterpret_cast<float4*>(y + (size_t)n * ldy + c);
    po[0] = make_float4(o[0], o[1], o[2], o[3]);
    po[1] = make_float4(o[4], o[5], o[6], o[7]);
    if (y16)   // operand copy for the tensor-core node GEMM that consumes y
      *reinterpret_cast<uint4*>(y16 + (size_t)n * ldy16 + c) = make_uint4(tc::pack_bf16(o[0], o[1]), tc::pack_bf16(o[2], o[3]),
                                                                         tc::pack_bf16(o[4], o[5]), tc::pack_bf16(o[6], o[7]));
  }
}

// g[k,:] = (ge[k,:] + gy[recv[k],:] / deg) * silu'(pre2[k,:]).  One warp per edge row PAIR: the receiver index and its
// degree are read once per row (lane 0, broadcast), every lane moves 16-byte pieces of both rows before it computes, and
// silu' is evaluated two elements per instruction in bf16x2 (the form the GEMM epilogues use) -- the kernel is a pure
// 2-pass stream (read pre2, write g) and was issue / latency bound in its one-thread-per-8-columns form (3.7 TB/s).
template <int LPR>     // lanes per row = H / 8
__global__ void edge_dpre2_bf16_kernel(int64_t E, const int32_t* __restrict__ recv, const int32_t* __restrict__ rowptr,
                                       const __nv_bfloat16* __restrict__ ge, const float* __restrict__ gy,
                                       const __nv_bfloat16* __restrict__ pre2, __nv_bfloat16* __restrict__ g) {
  constexpr int H = LPR * 8;
  constexpr int RPW = 32 / LPR;                      // rows per warp and pass
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int sub = lane / LPR, c = (lane % LPR) * 8;
#pragma unroll
  for (int p = 0; p < 2; ++p) {
    const int64_t k = (warp * 2 + p) * RPW + sub;
    if (k >= E) return;
    const int r = recv[k];
    const int deg = rowptr[r + 1] - rowptr[r];
    const float inv = 1.0f / (float)(deg < 1 ? 1 : deg);
    const uint4 z = __ldg(reinterpret_cast<const uint4*>(pre2 + (size_t)k * H + c));
    const float4 a0 = __ldg(reinterpret_cast<const float4*>(gy + (size_t)r * H + c));
    const float4 a1 = __ldg(reinterpret_cast<const float4*>(gy + (size_t)r * H + c + 4));
    float o[8] = {a0.x * inv, a0.y * inv, a0.z * inv, a0.w * inv, a1.x * inv, a1.y * inv, a1.z * inv, a1.w * inv};
    if (ge) {
      float v[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(ge + (size_t)k * H + c)), v);
#pragma unroll
      for (int q = 0; q < 8; ++q) o[q] += v[q];
    }
    const uint32_t zz[4] = {z.x, z.y, z.z, z.w};
    uint32_t w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t d = tc::bf2_dsilu_w(zz[q]);     // 2 silu'(z) - 1 for the pair
      const float h0 = 0.5f * o[2 * q], h1 = 0.5f * o[2 * q + 1];
      w[q] = tc::pack_bf16(fmaf(h0, bf16_lo(d), h0), fmaf(h1, bf16_hi(d), h1));
    }
    *reinterpret_cast<uint4*>(g + (size_t)k * H + c) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

__global__ void node_mean_scale_bf16_kernel(int64_t Nn, int H, const int32_t* __restrict__ rowptr,
                                            const float* __restrict__ src, __nv_bfloat16* __restrict__ out) {
  const int hq = H / 4;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Nn * hq) return;
  const int64_t n = i / hq;
  const int c = (int)(i % hq) * 4;
  const int deg = rowptr[n + 1] - rowptr[n];
  const float inv = 1.0f / (float)(deg < 1 ? 1 : deg);
  const float4 v = *reinterpret_cast<const float4*>(src + (size_t)n * H + c);
  uint2 o;
  o.x = tc::pack_bf16(v.x * inv, v.y * inv);
  o.y = tc::pack_bf16(v.z * inv, v.w * inv);
  *reinterpret_cast<uint2*>(out + (size_t)n * H + c) = o;
}

constexpr int CS_ROWS = 4096;

// partial[chunk][n] = sum of rows [chunk*CS_ROWS, ...) ; block = 32 column-octets x 8 row lanes
__global__ void colsum_bf16_partial_kernel(const __nv_bfloat16* __restrict__ A, int64_t lda, int64_t M, int N,
                                           float* __restrict__ partial) {
  __shared__ float sm[8][257];
  const int cg = threadIdx.x & 31, rl = threadIdx.x >> 5;   // column group (8 columns), row lane
  const int c = blockIdx.x * 256 + cg * 8;
  const int64_t r0 = (int64_t)blockIdx.y * CS_ROWS;
  const int64_t r1 = r0 + CS_ROWS < M ? r0 + CS_ROWS : M;
  float s[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = 0.0f;
  if (c < N)
    for (int64_t r = r0 + rl; r < r1; r += 8) {
      float v[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(A + (size_t)r * lda + c)), v);
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] += v[i];
    }
#pragma unroll
  for (int i = 0; i < 8; ++i) sm[rl][cg * 8 + i] = s[i];
  __syncthreads();
  const int col = threadIdx.x;
  if (blockIdx.x * 256 + col < N) {
    float t = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sm[i][col];
    partial[(size_t)blockIdx.y * N + blockIdx.x * 256 + col] = t;
  }
}

__global__ void colsum_final_kernel(const float* __restrict__ partial, int chunks, int N, float* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float s = 0.0f;
  for (int z = 0; z < chunks; ++z) s += partial[(size_t)z * N + n];
  out[n] = s;
}

template <bool SILU, bool MEAN>
int launch_segreduce(int64_t Nn, int W, const int32_t* ptr, const int32_t* list, const __nv_bfloat16* src, int64_t lds,
                     const float* base, int64_t ldb, float* y, int64_t ldy, cudaStream_t st, __nv_bfloat16* y16 = nullptr,
                     int64_t ldy16 = 0) {
  if (Nn <= 0) return 0;
  const unsigned blocks = (unsigned)((Nn * 32 + 255) / 256);
  if (W == 256) segreduce_bf16_kernel<32, SILU, MEAN><<<blocks, 256, 0, st>>>(Nn, ptr, list, src, lds, base, ldb, y, ldy, y16, ldy16);
  else if (W == 128) segreduce_bf16_kernel<16, SILU, MEAN><<<blocks, 256, 0, st>>>(Nn, ptr, list, src, lds, base, ldb, y, ldy, y16, ldy16);
  else if (W == 64) segreduce_bf16_kernel<8, SILU, MEAN><<<blocks, 256, 0, st>>>(Nn, ptr, list, src, lds, base, ldb, y, ldy, y16, ldy16);
  else return AETHER_E_UNSUPPORTED;
  AE_LAUNCH_CHECK();
  return 0;
}

}  // namespace

int launch_segmean_silu_bf16(int64_t Nn, int H, const int32_t* rowptr, const void* src, int64_t lds, const float* base,
                             int64_t ldb, float* y, int64_t ldy, cudaStream_t st, void* y16, int64_t ldy16) {
  return launch_segreduce<true, true>(Nn, H, rowptr, nullptr, (const __nv_bfloat16*)src, lds, base, ldb, y, ldy, st,
                                      (__nv_bfloat16*)y16, ldy16);
}

int launch_segsum_bf16(int64_t Nn, int W, const int32_t* ptr, const int32_t* list, const void* src, int64_t lds,
                       float* out, int64_t ldo, cudaStream_t st, void* out16, int64_t ldo16) {
  return launch_segreduce<false, false>(Nn, W, ptr, list, (const __nv_bfloat16*)src, lds, nullptr, 0, out, ldo, st,
                                        (__nv_bfloat16*)out16, ldo16);
}

int launch_edge_dpre2_bf16(int64_t E, int H, const int32_t* recv, const int32_t* rowptr, const void* ge, const float* gy,
                           const void* pre2, void* g, cudaStream_t st) {
  if (E <= 0) return 0;
  if (H != 64 && H != 128 && H != 256) return AETHER_E_BADARG;
  const int rpw = 2 * (32 / (H / 8));                // rows per warp
  const int64_t warps = (E + rpw - 1) / rpw;
  const unsigned blocks = (unsigned)((warps * 32 + 255) / 256);
  auto a = (const __nv_bfloat16*)ge; auto z = (const __nv_bfloat16*)pre2; auto o = (__nv_bfloat16*)g;
  if (H == 256) edge_dpre2_bf16_kernel<32><<<blocks, 256, 0, st>>>(E, recv, rowptr, a, gy, z, o);
  else if (H == 128) edge_dpre2_bf16_kernel<16><<<blocks, 256, 0, st>>>(E, recv, rowptr, a, gy, z, o);
  else edge_dpre2_bf16_kernel<8><<<blocks, 256, 0, st>>>(E, recv, rowptr, a, gy, z, o);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_node_mean_scale_bf16(int64_t Nn, int H, const int32_t* rowptr, const float* src, void* out16, cudaStream_t st) {
  if (Nn <= 0) return 0;
  const int64_t tot = Nn * (H / 4);
  node_mean_scale_bf16_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(Nn, H, rowptr, src, (__nv_bfloat16*)out16);
  AE_LAUNCH_CHECK();
  return 0;
}

size_t colsum_bf16_partial_floats(int64_t M, int N) { return (size_t)ceil_div64(M > 0 ? M : 1, CS_ROWS) * N; }

int colsum_bf16(const void* A, int64_t lda, int64_t M, int N, float* out, float* partial, size_t partial_floats,
                cudaStream_t st) {
  if (N <= 0) return 0;
  if (N % 8) return AETHER_E_BADARG;
  const int chunks = (int)ceil_div64(M > 0 ? M : 1, CS_ROWS);
  if (partial_floats < (size_t)chunks * N) return AETHER_E_WORKSPACE;
  dim3 grid((N + 255) / 256, chunks);
  colsum_bf16_partial_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)A, lda, M, N, partial);
  AE_LAUNCH_CHECK();
  colsum_final_kernel<<<(N + 255) / 256, 256, 0, st>>>(partial, chunks, N, out);
  AE_LAUNCH_CHECK();
  return 0;
}

}  // namespace aether
