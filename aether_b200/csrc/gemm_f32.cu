// Generic fp32 GEMM family on CUDA cores with fused load transforms and epilogues.
//
// Used for every Linear of the step in the fp32 path (nn/state2state/locs/locs.py:206-225,
// 160-168; nn/state2state/aether.py:113-119) and for their backward (data and weight
// gradients).  The bf16 tcgen05 kernels in edge_mlp_tc.cu replace the edge-level calls when
// AETHER_F_BF16_EDGE_MLP is set; node-level GEMMs (Nn rows) always run here.
//
// Tile: BM=128 x BN (64|128) x BK=16, 256 threads, 8 x (BN/16) register tile per thread,
// shared-memory double buffering with register prefetch of the next K slab.
#include "common.cuh"

namespace aether {

namespace {

constexpr int BM = 128;
constexpr int BK = 16;
constexpr int NT = 256;
constexpr int APAD = 4;

struct TileCoord { int64_t m0; int n0; int64_t k_begin; int64_t k_end; };

// Element of the "A side" (rows of C) for reduction index k.
template <int LAYOUT>
__device__ __forceinline__ float load_a(const GemmArgs& g, int64_t m, int64_t k) {
  if (m >= g.M || k >= g.K) return 0.0f;
  if (LAYOUT == GEMM_TN) return g.A[(size_t)k * g.lda + m];
  float v = g.A[(size_t)m * g.lda + k];
  if (LAYOUT == GEMM_NT && g.tf != TF_NONE) {
    v = silu_f(v);
    if (g.tf == TF_SILU_MASK) v *= g.tf_mask[(size_t)m * g.lda + k];
  }
  return v;
}

// Element of the "B side" (columns of C) for reduction index k.
template <int LAYOUT>
__device__ __forceinline__ float load_b(const GemmArgs& g, int64_t k, int n) {
  if (n >= g.N || k >= g.K) return 0.0f;
  if (LAYOUT == GEMM_NT) return g.B[(size_t)n * g.ldb + k];
  float v = g.B[(size_t)k * g.ldb + n];
  if (LAYOUT == GEMM_TN && g.tf != TF_NONE) {
    v = silu_f(v);
    if (g.tf == TF_SILU_MASK) v *= g.tf_mask[(size_t)k * g.ldb + n];
  }
  return v;
}

template <int LAYOUT, int BN>
__global__ void __launch_bounds__(NT, 2) gemm_kernel(GemmArgs g, int64_t k_chunk) {
  constexpr int TN_ = BN / 16;          // 4 or 8 columns per thread
  constexpr int A_PER = BM * BK / NT;   // 8
  constexpr int B_PER = BN * BK / NT;   // 4 or 8
  __shared__ __align__(16) float As[2][BK][BM + APAD];
  __shared__ __align__(16) float Bs[2][BK][BN + APAD];

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  int64_t k_begin = 0, k_end = g.K;
  if (LAYOUT == GEMM_TN) {
    k_begin = (int64_t)blockIdx.z * k_chunk;
    k_end = k_begin + k_chunk < g.K ? k_begin + k_chunk : g.K;
  }

  // Which operand is contiguous along k decides the global->shared mapping.
  //   k-contiguous rows (NT: A and B, NN: A): thread reads (row = t/16 + 16 i, k = t%16)
  //   mn-contiguous rows (TN: A and B, NN: B): thread reads (k = t/W + ..., col = t%W)
  float ra[A_PER], rb[B_PER];

  auto fetch = [&](int64_t k0) {
#pragma unroll
    for (int i = 0; i < A_PER; ++i) {
      if (LAYOUT == GEMM_TN) {
        int col = tid & (BM - 1), kk = (tid >> 7) + 2 * i;
        ra[i] = load_a<LAYOUT>(g, m0 + col, (k0 + kk < k_end) ? k0 + kk : g.K);
      } else {
        int row = (tid >> 4) + 16 * i, kk = tid & 15;
        ra[i] = load_a<LAYOUT>(g, m0 + row, (k0 + kk < k_end) ? k0 + kk : g.K);
      }
    }
#pragma unroll
    for (int i = 0; i < B_PER; ++i) {
      if (LAYOUT == GEMM_NT) {
        int row = (tid >> 4) + 16 * i, kk = tid & 15;
        rb[i] = load_b<LAYOUT>(g, (k0 + kk < k_end) ? k0 + kk : g.K, n0 + row);
      } else {
        int col = tid & (BN - 1), kk = tid / BN + (NT / BN) * i;
        rb[i] = load_b<LAYOUT>(g, (k0 + kk < k_end) ? k0 + kk : g.K, n0 + col);
      }
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int i = 0; i < A_PER; ++i) {
      if (LAYOUT == GEMM_TN) {
        int col = tid & (BM - 1), kk = (tid >> 7) + 2 * i;
        As[buf][kk][col] = ra[i];
      } else {
        int row = (tid >> 4) + 16 * i, kk = tid & 15;
        As[buf][kk][row] = ra[i];
      }
    }
#pragma unroll
    for (int i = 0; i < B_PER; ++i) {
      if (LAYOUT == GEMM_NT) {
        int row = (tid >> 4) + 16 * i, kk = tid & 15;
        Bs[buf][kk][row] = rb[i];
      } else {
        int col = tid & (BN - 1), kk = tid / BN + (NT / BN) * i;
        Bs[buf][kk][col] = rb[i];
      }
    }
  };

  float acc[8][TN_];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < TN_; ++j) acc[i][j] = 0.0f;

  int buf = 0;
  if (k_begin < k_end) {
    fetch(k_begin);
    stash(0);
  }
  __syncthreads();
  for (int64_t k0 = k_begin; k0 < k_end; k0 += BK) {
    const bool more = k0 + BK < k_end;
    if (more) fetch(k0 + BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[8], b[TN_];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 8 + 4]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w;
      a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w;
      if (TN_ == 8) {
        const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
        b[TN_ - 4] = b1.x; b[TN_ - 3] = b1.y; b[TN_ - 2] = b1.z; b[TN_ - 1] = b1.w;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < TN_; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (more) {
      stash(buf ^ 1);
      __syncthreads();
      buf ^= 1;
    }
  }

  // ---- epilogue ----
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t m = m0 + ty * 8 + i;
    if (m >= g.M) continue;
    int ia = 0, ib = 0;
    if (LAYOUT != GEMM_TN && g.gat) { ia = g.idx_a[m]; ib = g.idx_b[m]; }
#pragma unroll
    for (int j = 0; j < TN_; ++j) {
      const int n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      if (n >= g.N) continue;
      float v = acc[i][j];
      if (LAYOUT == GEMM_TN) {
        g.partial[((size_t)blockIdx.z * g.M + m) * g.N + n] = v;
        continue;
      }
      if (g.bias) v += g.bias[n];
      if (g.gat) v += g.gat[(size_t)ia * g.ldg + n] + g.gat[(int64_t)ib * g.ldg + g.goff_b + n];
      if (g.res) v += g.res[(size_t)m * g.ldr + n];
      if (g.dsilu) v *= dsilu_f(g.dsilu[(size_t)m * g.ldd + n]);
      if (g.emask) v *= g.emask[(size_t)m * g.ldm + n];
      float* c = g.C + (size_t)m * g.ldc + n;
      if (g.accumulate) v += *c;
      *c = v;
    }
  }
}

// C[i] (=|+=) sum_z partial[z][i], z ascending: deterministic.
__global__ void reduce_partials_kernel(const float* __restrict__ partial, int splits, size_t mn, int N, float* C,
                                       int64_t ldc, int accumulate) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= mn) return;
  float s = 0.0f;
  for (int z = 0; z < splits; ++z) s += partial[(size_t)z * mn + i];
  size_t m = i / N, n = i % N;
  float* c = C + m * ldc + n;
  *c = accumulate ? *c + s : s;
}

constexpr int COLSUM_ROWS = 2048;

__global__ void colsum_partial_kernel(const float* __restrict__ A, int64_t lda, int64_t M, int N, float* partial) {
  // block (x: column group of 32, y: row chunk); 32 x 8 threads
  __shared__ float sm[8][33];
  const int n = blockIdx.x * 32 + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * COLSUM_ROWS;
  const int64_t r1 = r0 + COLSUM_ROWS < M ? r0 + COLSUM_ROWS : M;
  float s = 0.0f;
  if (n < N)
    for (int64_t r = r0 + threadIdx.y; r < r1; r += 8) s += A[(size_t)r * lda + n];
  sm[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && n < N) {
    float t = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sm[i][threadIdx.x];
    partial[(size_t)blockIdx.y * N + n] = t;
  }
}

int tn_splits(int64_t M, int N, int64_t K) {
  int64_t tiles = ceil_div64(M, BM) * ceil_div64(N, 64);
  int64_t want = ceil_div64(148 * 3, tiles);
  int64_t max_by_k = ceil_div64(K, 4 * BK);   // at least 64 reduction rows per split
  int64_t s = want < max_by_k ? want : max_by_k;
  if (s < 1) s = 1;
  if (s > 512) s = 512;
  return (int)s;
}

}  // namespace

size_t gemm_tn_partial_floats(int64_t M, int N, int64_t K) { return (size_t)tn_splits(M, N, K) * M * N; }

int gemm_f32(int layout, const GemmArgs& a, cudaStream_t st) {
  if (a.M <= 0 || a.N <= 0) return 0;
  {
    // Linears with one thin side (2..12 wide next to the hidden size) have kernels of their own (thin.cu)
    const int rc = gemm_f32_thin(layout, a, st);
    if (rc != AETHER_E_UNSUPPORTED) return rc;
  }
  const bool wide = a.N > 64 && a.M >= 4096;
  dim3 grid((unsigned)ceil_div64(a.M, BM), (unsigned)ceil_div64(a.N, wide ? 128 : 64), 1);
  if (layout == GEMM_NT) {
    if (wide) gemm_kernel<GEMM_NT, 128><<<grid, NT, 0, st>>>(a, 0);
    else gemm_kernel<GEMM_NT, 64><<<grid, NT, 0, st>>>(a, 0);
  } else if (layout == GEMM_NN) {
    if (wide) gemm_kernel<GEMM_NN, 128><<<grid, NT, 0, st>>>(a, 0);
    else gemm_kernel<GEMM_NN, 64><<<grid, NT, 0, st>>>(a, 0);
  } else {
    const int splits = tn_splits(a.M, a.N, a.K);
    if (!a.partial || a.partial_floats < (size_t)splits * a.M * a.N) return AETHER_E_WORKSPACE;
    int64_t chunk = ceil_div64(ceil_div64(a.K > 0 ? a.K : 1, splits), BK) * BK;
    grid = dim3((unsigned)ceil_div64(a.M, BM), (unsigned)ceil_div64(a.N, 64), (unsigned)splits);
    gemm_kernel<GEMM_TN, 64><<<grid, NT, 0, st>>>(a, chunk);
    AE_LAUNCH_CHECK();
    size_t mn = (size_t)a.M * a.N;
    reduce_partials_kernel<<<(unsigned)((mn + 255) / 256), 256, 0, st>>>(a.partial, splits, mn, a.N, a.C, a.ldc,
                                                                         a.accumulate);
  }
  AE_LAUNCH_CHECK();
  return 0;
}

size_t colsum_partial_floats(int64_t M, int N) { return (size_t)ceil_div64(M > 0 ? M : 1, COLSUM_ROWS) * N; }

int colsum_f32(const float* A, int64_t lda, int64_t M, int N, float* out, float* partial, size_t partial_floats,
               cudaStream_t st) {
  if (N <= 0) return 0;
  const int chunks = (int)ceil_div64(M > 0 ? M : 1, COLSUM_ROWS);
  if (partial_floats < (size_t)chunks * N) return AETHER_E_WORKSPACE;
  dim3 grid((N + 31) / 32, chunks);
  colsum_partial_kernel<<<grid, dim3(32, 8), 0, st>>>(A, lda, M, N, partial);
  AE_LAUNCH_CHECK();
  reduce_partials_kernel<<<(N + 255) / 256, 256, 0, st>>>(partial, chunks, (size_t)N, N, out, N, 0);
  AE_LAUNCH_CHECK();
  return 0;
}

}  // namespace aether
