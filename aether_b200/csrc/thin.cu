// fp32 Linears with one thin side, node level (same contract as gemm_f32.cu's GemmArgs; gemm_f32() dispatches here).
//
// Three Linears of the step have a dimension of 2..9 next to the hidden size: the residual Linear of layer 1
// (3D -> H, locs.py:214-218), the last Linear of the out MLP (H -> D, locs.py:160-168) and, in the fp32 path, nothing
// else.  Forward, data gradient and weight gradient of each are GEMMs in which the 128 x 64 x 16 tiles of the generic
// kernel are > 90 % padding: 140-450 us per launch at 102 400 rows, where the traffic ([rows, H] fp32 once) is 16 us.
//   expand : C[m, :] = epi(sum_{k < K <= 16} a[m, k] w[k, :])       wide output, thin reduction     (NT / NN)
//   reduce : C[m, n < N <= 12] = epi(sum_k tf(a[m, k]) w[k, n])     thin output, wide reduction     (NT / NN)
//   tn     : C = A^T tf(B) over all rows, one side <= 12 wide       weight gradients, deterministic two-stage sum
// Every element goes through the same epilogue order as gemm_kernel (bias, res, silu', mask, accumulate).
#include "common.cuh"

namespace aether {
namespace {

constexpr int kThinK = 16;      // expand: reduction length held in registers
constexpr int kThinN = 12;      // reduce / tn: width of the thin side held in registers

__device__ __forceinline__ float epilogue(const GemmArgs& g, int64_t m, int n, float v) {
  if (g.bias) v += g.bias[n];
  if (g.res) v += g.res[(size_t)m * g.ldr + n];
  if (g.dsilu) v *= dsilu_f(g.dsilu[(size_t)m * g.ldd + n]);
  if (g.emask) v *= g.emask[(size_t)m * g.ldm + n];
  return v;
}

// ---- expand: block = 256 threads, thread = 4 adjacent output columns, rows strided over the block's row range ----------
template <int LAYOUT>
__global__ void __launch_bounds__(256) thin_expand_kernel(GemmArgs g, int rows_per_block) {
  extern __shared__ float Ws[];                       // [K][N]
  const int K = (int)g.K, N = g.N;
  for (int i = threadIdx.x; i < K * N; i += 256) {
    const int k = i / N, n = i - k * N;
    Ws[i] = LAYOUT == GEMM_NT ? g.B[(size_t)n * g.ldb + k] : g.B[(size_t)k * g.ldb + n];
  }
  __syncthreads();
  const int tpr = N / 4;                              // threads per row
  const int rstep = 256 / tpr > 0 ? 256 / tpr : 1;    // rows in flight (N <= 1024)
  const int c = (threadIdx.x % tpr) * 4;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = r0 + rows_per_block < g.M ? r0 + rows_per_block : g.M;
  if (threadIdx.x >= tpr * rstep) return;
  for (int64_t m = r0 + threadIdx.x / tpr; m < r1; m += rstep) {
    float a[kThinK];
#pragma unroll
    for (int k = 0; k < kThinK; ++k) {
      float v = 0.0f;
      if (k < K) {
        v = g.A[(size_t)m * g.lda + k];
        if (LAYOUT == GEMM_NT && g.tf != TF_NONE) {
          v = silu_f(v);
          if (g.tf == TF_SILU_MASK) v *= g.tf_mask[(size_t)m * g.lda + k];
        }
      }
      a[k] = v;
    }
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < kThinK; ++k)
      if (k < K) {
        const float4 w = *reinterpret_cast<const float4*>(Ws + k * N + c);
        acc[0] = fmaf(a[k], w.x, acc[0]); acc[1] = fmaf(a[k], w.y, acc[1]);
        acc[2] = fmaf(a[k], w.z, acc[2]); acc[3] = fmaf(a[k], w.w, acc[3]);
      }
    float4* dst = reinterpret_cast<float4*>(g.C + (size_t)m * g.ldc + c);
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = epilogue(g, m, c + j, acc[j]);
    if (g.accumulate) { const float4 p = *dst; o[0] += p.x; o[1] += p.y; o[2] += p.z; o[3] += p.w; }
    *dst = make_float4(o[0], o[1], o[2], o[3]);
  }
}

// ---- reduce: one warp per row, lanes stride the reduction with 128-bit loads, N accumulators, shuffle tree -------------
template <int LAYOUT>
__global__ void __launch_bounds__(256) thin_reduce_kernel(GemmArgs g) {
  extern __shared__ __align__(16) float Ws[];         // [N][K]
  const int K = (int)g.K, N = g.N;
  for (int i = threadIdx.x; i < K * N; i += 256) {
    const int n = i / K, k = i - n * K;
    Ws[i] = LAYOUT == GEMM_NT ? g.B[(size_t)n * g.ldb + k] : g.B[(size_t)k * g.ldb + n];
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5), nwarps = (int64_t)gridDim.x * 8;
  for (int64_t m = warp; m < g.M; m += nwarps) {
    float acc[kThinN];
#pragma unroll
    for (int n = 0; n < kThinN; ++n) acc[n] = 0.0f;
    const float* arow = g.A + (size_t)m * g.lda;
    for (int k = lane * 4; k < K; k += 128) {
      float4 a = *reinterpret_cast<const float4*>(arow + k);
      if (LAYOUT == GEMM_NT && g.tf != TF_NONE) {
        a.x = silu_f(a.x); a.y = silu_f(a.y); a.z = silu_f(a.z); a.w = silu_f(a.w);
        if (g.tf == TF_SILU_MASK) {
          const float4 mk = *reinterpret_cast<const float4*>(g.tf_mask + (size_t)m * g.lda + k);
          a.x *= mk.x; a.y *= mk.y; a.z *= mk.z; a.w *= mk.w;
        }
      }
#pragma unroll
      for (int n = 0; n < kThinN; ++n)
        if (n < N) {
          const float4 w = *reinterpret_cast<const float4*>(Ws + n * K + k);
          acc[n] = fmaf(a.x, w.x, acc[n]); acc[n] = fmaf(a.y, w.y, acc[n]);
          acc[n] = fmaf(a.z, w.z, acc[n]); acc[n] = fmaf(a.w, w.w, acc[n]);
        }
    }
    float mine = 0.0f;
#pragma unroll
    for (int n = 0; n < kThinN; ++n)
      if (n < N) {
        float v = acc[n];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == n) mine = v;
      }
    if (lane < N) {
      float v = epilogue(g, m, lane, mine);
      float* c = g.C + (size_t)m * g.ldc + lane;
      if (g.accumulate) v += *c;
      *c = v;
    }
  }
}

// ---- tn: C[M, N] = sum_k A[k, m] tf(B[k, n]) with one of M, N thin.  Block = 256 threads = 256 columns of the wide
//      side, grid.y = row chunks; every thread keeps the thin side's T accumulators and walks its chunk's rows (wide
//      element coalesced over the block, thin elements broadcast).  Chunk partials -> ordered sum: deterministic. ------
template <bool THIN_A>
__global__ void __launch_bounds__(256) thin_tn_kernel(GemmArgs g, int64_t k_chunk) {
  const int W = THIN_A ? g.N : (int)g.M, T = THIN_A ? (int)g.M : g.N;
  const int j = blockIdx.x * 256 + threadIdx.x;
  const int64_t k0 = (int64_t)blockIdx.y * k_chunk, k1 = k0 + k_chunk < g.K ? k0 + k_chunk : g.K;
  float acc[kThinN];
#pragma unroll
  for (int t = 0; t < kThinN; ++t) acc[t] = 0.0f;
  const bool on = j < W;
  const int jj = on ? j : 0;
  auto b_tf = [&](int64_t k, int n) {
    float v = g.B[(size_t)k * g.ldb + n];
    if (g.tf != TF_NONE) {
      v = silu_f(v);
      if (g.tf == TF_SILU_MASK) v *= g.tf_mask[(size_t)k * g.ldb + n];
    }
    return v;
  };
#pragma unroll 4
  for (int64_t k = k0; k < k1; ++k) {
    const float wide = THIN_A ? b_tf(k, jj) : g.A[(size_t)k * g.lda + jj];
#pragma unroll
    for (int t = 0; t < kThinN; ++t)
      if (t < T) {
        const float thin = THIN_A ? g.A[(size_t)k * g.lda + t] : b_tf(k, t);
        acc[t] = fmaf(wide, thin, acc[t]);
      }
  }
  if (!on) return;
#pragma unroll
  for (int t = 0; t < kThinN; ++t)
    if (t < T) {
      const int m = THIN_A ? t : j, n = THIN_A ? j : t;
      g.partial[((size_t)blockIdx.y * g.M + m) * g.N + n] = acc[t];
    }
}

__global__ void thin_reduce_partials_kernel(const float* __restrict__ partial, int splits, size_t mn, int N, float* C,
                                            int64_t ldc, int accumulate) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= mn) return;
  float s = 0.0f;
  for (int z = 0; z < splits; ++z) s += partial[(size_t)z * mn + i];
  float* c = C + (i / N) * ldc + i % N;
  *c = accumulate ? *c + s : s;
}

inline bool al16(const void* p) { return ((uintptr_t)p & 15) == 0; }

}  // namespace

// Returns AETHER_E_UNSUPPORTED when the shape is not a thin one (the caller then runs the generic kernel).
int gemm_f32_thin(int layout, const GemmArgs& a, cudaStream_t st) {
  if (a.gat || a.M <= 0 || a.N <= 0 || a.K <= 0) return AETHER_E_UNSUPPORTED;
  if (layout == GEMM_NT || layout == GEMM_NN) {
    // (the epilogue reads its operands element-wise; only C and, in reduce, the A rows are accessed as float4)
    if (a.K <= kThinK && a.N >= 32 && a.N <= 1024 && a.N % 4 == 0 && a.ldc % 4 == 0 && al16(a.C) &&
        (size_t)a.K * a.N * sizeof(float) <= 48 * 1024) {
      const int rows_per_block = 64;
      const size_t smem = (size_t)a.K * a.N * sizeof(float);
      const unsigned grid = (unsigned)ceil_div64(a.M, rows_per_block);
      if (layout == GEMM_NT) thin_expand_kernel<GEMM_NT><<<grid, 256, smem, st>>>(a, rows_per_block);
      else thin_expand_kernel<GEMM_NN><<<grid, 256, smem, st>>>(a, rows_per_block);
      AE_LAUNCH_CHECK();
      return 0;
    }
    if (a.N <= kThinN && a.K >= 32 && a.K <= 1024 && a.K % 4 == 0 && a.lda % 4 == 0 && al16(a.A) &&
        (a.tf != TF_SILU_MASK || al16(a.tf_mask))) {
      const size_t smem = (size_t)a.K * a.N * sizeof(float);
      int64_t blocks = ceil_div64(a.M, 8);
      if (blocks > 148 * 8) blocks = 148 * 8;
      if (layout == GEMM_NT) thin_reduce_kernel<GEMM_NT><<<(unsigned)blocks, 256, smem, st>>>(a);
      else thin_reduce_kernel<GEMM_NN><<<(unsigned)blocks, 256, smem, st>>>(a);
      AE_LAUNCH_CHECK();
      return 0;
    }
    return AETHER_E_UNSUPPORTED;
  }
  // TN
  const bool thin_a = a.M <= kThinN && a.N >= 32, thin_b = a.N <= kThinN && a.M >= 32;
  if (!thin_a && !thin_b) return AETHER_E_UNSUPPORTED;
  const size_t mn = (size_t)a.M * a.N;
  if (!a.partial || a.partial_floats < mn) return AETHER_E_UNSUPPORTED;
  // row chunks: about four blocks per SM over the whole grid, at least 32 rows each (a thread walks its chunk serially)
  const int64_t col_blocks = ((thin_a ? a.N : a.M) + 255) / 256;
  int64_t want = ceil_div64(148 * 4, col_blocks);
  int64_t splits = ceil_div64(a.K, 32) < want ? ceil_div64(a.K, 32) : want;
  if (splits < 1) splits = 1;
  if ((size_t)splits > a.partial_floats / mn) splits = (int64_t)(a.partial_floats / mn);
  const int64_t chunk = ceil_div64(a.K, splits);
  splits = ceil_div64(a.K, chunk);
  const int W = thin_a ? a.N : (int)a.M;
  dim3 grid((unsigned)((W + 255) / 256), (unsigned)splits);
  if (thin_a) thin_tn_kernel<true><<<grid, 256, 0, st>>>(a, chunk);
  else thin_tn_kernel<false><<<grid, 256, 0, st>>>(a, chunk);
  AE_LAUNCH_CHECK();
  thin_reduce_partials_kernel<<<(unsigned)((mn + 255) / 256), 256, 0, st>>>(a.partial, (int)splits, mn, a.N, a.C, a.ldc,
                                                                            a.accumulate);
  AE_LAUNCH_CHECK();
  return 0;
}

}  // namespace aether
