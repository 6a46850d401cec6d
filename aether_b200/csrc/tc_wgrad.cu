// Weight-gradient GEMM on tcgen05:   dW[MG, NA] = sum over rows e of  G[e, :]^T  tf(ACT[e, :])
//
// (backward of every edge-level Linear of GNNLayer.message_fn, nn/state2state/locs/locs.py:206-212:
// G is the gradient at the Linear's output, ACT the saved pre-activation of its input, SiLU applied on
// load.)  The reduction runs over edges, so both operands are "MN-major" for the tensor core: a TMA box
// of 64 rows x 64 bf16 columns with the 128-byte swizzle is exactly the UMMA MN-major layout, one 8 KB
// slab per 64 columns.  Each CTA owns a contiguous range of 64-row stages, accumulates its partial dW in
// TMEM (MG/128 accumulators of NA columns: all 512 columns at MG = NA = 256) and writes it once; a
// second kernel sums the per-CTA partials in a fixed order, so the result is deterministic.
//
//   warp 4     TMA producer (MG/64 + NA/64 boxes per stage)
//   warps 6-9  in-place SiLU on the ACT slabs
//   warp 5     MMA issuer: per stage 4 K-steps x MG/128 instructions of 128 x NA x 16
//   warps 0-3  final TMEM -> partial buffer
#include "tc_gemm.cuh"

namespace aether {

namespace {

using namespace tc;

constexpr int kThreads = 320;
constexpr int kStageRows = 64;
constexpr int kSlab = kStageRows * 128;           // 64 rows x 64 bf16 = 8 KB
constexpr int kSmemMax = 232448;

__host__ __device__ constexpr int stage_bytes(int MG, int NA) { return (MG / 64 + NA / 64) * kSlab; }
__host__ __device__ constexpr int wg_stages(int MG, int NA) {
  int s = (kSmemMax - 1024 - 1024) / stage_bytes(MG, NA);
  return s > 6 ? 6 : s;
}
__host__ __device__ constexpr int wg_smem(int MG, int NA) { return 1024 + wg_stages(MG, NA) * stage_bytes(MG, NA) + 1024; }
__host__ __device__ constexpr int wg_mtiles(int MG) { return MG >= 128 ? MG / 128 : 1; }
__host__ __device__ constexpr int wg_tmem_cols(int MG, int NA) {
  int c = wg_mtiles(MG) * NA;
  return c < 32 ? 32 : c;
}

template <int MG, int NA, int TF>
__global__ void __launch_bounds__(kThreads, 1)
tc_wgrad_kernel(const __grid_constant__ CUtensorMap tmG, const __grid_constant__ CUtensorMap tmA, int64_t n_stages,
                float* __restrict__ partial, float* __restrict__ colsum_partial) {
  static_assert(MG % 64 == 0 && NA % 64 == 0 && MG <= 256 && NA <= 256, "unsupported shape");
  constexpr int S = wg_stages(MG, NA);
  constexpr int MT = wg_mtiles(MG);
  constexpr int GA = MG / 64, GB = NA / 64;       // slabs per stage
  constexpr int STAGE = stage_bytes(MG, NA);
  constexpr uint32_t IDESC = make_idesc(128, NA, MAJOR_MN, MAJOR_MN);
  // MG == 64: rows 64..127 of the M=128 instruction re-read slab 0 (LBO = 0); those lanes are ignored
  constexpr uint32_t LBO_A = MG >= 128 ? kSlab : 0;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sMisc = base + S * STAGE;
  const uint32_t bar_raw = sMisc, bar_rdy = sMisc + 64, bar_free = sMisc + 128, bar_done = sMisc + 192;
  const uint32_t tmem_slot = sMisc + 208;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // contiguous range of stages of this CTA
  const int64_t s_begin = n_stages * blockIdx.x / gridDim.x;
  const int64_t s_end = n_stages * (blockIdx.x + 1) / gridDim.x;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(bar_raw + 8 * s, 1);
      mbar_init(bar_rdy + 8 * s, 128);
      mbar_init(bar_free + 8 * s, colsum_partial ? 5 : 1);   // MMA commit (+ the four column-sum warps)
    }
    mbar_init(bar_done, 1);
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc(tmem_slot, wg_tmem_cols(MG, NA));
  if (warp == 4 && lane == 0) { tma_prefetch_desc(&tmG); tma_prefetch_desc(&tmA); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 4) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t st = s_begin; st < s_end; ++st, ++it) {
        const uint32_t s = it % S, ph = (it / S) & 1;
        mbar_wait(bar_free + 8 * s, ph ^ 1);
        mbar_arrive_expect_tx(bar_raw + 8 * s, STAGE);
        const uint32_t dst = base + s * STAGE;
        const int r0 = (int)(st * kStageRows);
#pragma unroll
        for (int i = 0; i < GA; ++i) tma_load_2d(dst + i * kSlab, &tmG, i * 64, r0, bar_raw + 8 * s);
#pragma unroll
        for (int i = 0; i < GB; ++i) tma_load_2d(dst + (GA + i) * kSlab, &tmA, i * 64, r0, bar_raw + 8 * s);
      }
    }
  } else if (warp == 5) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t st = s_begin; st < s_end; ++st, ++it) {
        const uint32_t s = it % S, ph = (it / S) & 1;
        mbar_wait((TF != TC_TF_NONE ? bar_rdy : bar_raw) + 8 * s, ph);
        tc_fence_after();
        const uint32_t sa = base + s * STAGE, sb = sa + GA * kSlab;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t bd = make_sdesc(sb + k * 2048, kSlab, 1024);
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            const uint64_t ad = make_sdesc(sa + mt * 2 * kSlab + k * 2048, LBO_A, 1024);
            umma_bf16(tmem_base + mt * NA, ad, bd, IDESC, (it | (uint32_t)k) != 0);
          }
        }
        umma_commit(bar_free + 8 * s);
      }
      umma_commit(bar_done);
    }
  } else if (warp >= 6) {
    if (TF != TC_TF_NONE) {
      const int tt = threadIdx.x - 192;
      uint32_t it = 0;
      for (int64_t st = s_begin; st < s_end; ++st, ++it) {
        const uint32_t s = it % S, ph = (it / S) & 1;
        mbar_wait(bar_raw + 8 * s, ph);
        const uint32_t a0 = base + s * STAGE + GA * kSlab + tt * 16;
#pragma unroll
        for (int i = 0; i < GB * 4; ++i) {
          uint32_t u[4];
          const uint32_t addr = a0 + i * 2048;
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]) : "r"(addr));
#pragma unroll
          for (int q = 0; q < 4; ++q) u[q] = pack_bf16(silu_fast(bf16_lo(u[q])), silu_fast(bf16_hi(u[q])));
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[3]) : "memory");
        }
        fence_proxy_async_smem();
        mbar_arrive(bar_rdy + 8 * s);
      }
    }
  } else {
    // warps 0..3.  While the pipeline runs they are otherwise idle: optionally accumulate the column sums
    // of G (the bias gradient of the same Linear) from the landed slabs.  Thread t owns columns 2t, 2t+1:
    // slab t/32, 16-byte chunk (t%32)/4 (XOR-swizzled by the row), conflict-free 4-byte loads.
    if (colsum_partial) {
      const int t = threadIdx.x;
      float c0 = 0.0f, c1 = 0.0f;
      if (2 * t < MG) {
        const uint32_t slab_off = (uint32_t)(t >> 5) * kSlab + (uint32_t)(t & 3) * 4;
        const uint32_t chunk = (uint32_t)(t & 31) >> 2;
        uint32_t it = 0;
        for (int64_t st = s_begin; st < s_end; ++st, ++it) {
          const uint32_t s = it % S, ph = (it / S) & 1;
          mbar_wait(bar_raw + 8 * s, ph);
          const uint32_t a0 = base + s * STAGE + slab_off;
#pragma unroll 16
          for (int k = 0; k < kStageRows; ++k) {
            uint32_t v;
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(a0 + k * 128 + ((chunk ^ (uint32_t)(k & 7)) << 4)));
            c0 += bf16_lo(v);
            c1 += bf16_hi(v);
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_free + 8 * s);
        }
        colsum_partial[(size_t)blockIdx.x * MG + 2 * t] = c0;
        colsum_partial[(size_t)blockIdx.x * MG + 2 * t + 1] = c1;
      } else {
        // MG < 256: the upper warps only release the slots
        uint32_t it = 0;
        for (int64_t st = s_begin; st < s_end; ++st, ++it) {
          const uint32_t s = it % S, ph = (it / S) & 1;
          mbar_wait(bar_raw + 8 * s, ph);
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_free + 8 * s);
        }
      }
    }
    // then wait for the last MMA and write this CTA's partial dW
    float* out = partial + (size_t)blockIdx.x * MG * NA;
    if (s_end > s_begin) {
      mbar_wait(bar_done, 0);
      tc_fence_after();
    }
#pragma unroll 1
    for (int mt = 0; mt < MT; ++mt) {
      const int m = mt * 128 + warp * 32 + lane;
#pragma unroll 1
      for (int j = 0; j < NA / 32; ++j) {
        uint32_t r[32];
        if (s_end > s_begin) {
          tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + mt * NA + j * 32, r);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = 0u;
        }
        if (m < MG) {
          float4* po = reinterpret_cast<float4*>(out + (size_t)m * NA + j * 32);
#pragma unroll
          for (int q = 0; q < 8; ++q)
            po[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]), __uint_as_float(r[4 * q + 2]),
                                __uint_as_float(r[4 * q + 3]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, wg_tmem_cols(MG, NA));
  }
}

// dW[m*ld + n] = sum_b partial[b][m][n (+ 32 + n if fold32)], b ascending
__global__ void wgrad_reduce_kernel(const float* __restrict__ partial, int parts, int MG, int NA, int n_valid, int fold32,
                                    float* __restrict__ dW, int64_t ld) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= MG * n_valid) return;
  const int m = i / n_valid, n = i % n_valid;
  float s = 0.0f;
  const size_t mn = (size_t)MG * NA;
  for (int b = 0; b < parts; ++b) {
    const float* p = partial + (size_t)b * mn + (size_t)m * NA;
    s += fold32 ? p[n] + p[32 + n] : p[n];
  }
  dW[(size_t)m * ld + n] = s;
}

__global__ void colsum_reduce_kernel(const float* __restrict__ partial, int parts, int MG, float* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= MG) return;
  float s = 0.0f;
  for (int b = 0; b < parts; ++b) s += partial[(size_t)b * MG + n];
  out[n] = s;
}

int wg_grid(int64_t n_stages) {
  const int sms = tc_sm_count();
  return (int)(n_stages < sms ? (n_stages > 0 ? n_stages : 1) : sms);
}

template <int MG, int NA>
int launch_wg(const CUtensorMap& tg, const CUtensorMap& ta, int tf, int64_t n_stages, int grid, float* partial,
              float* cpart, cudaStream_t st) {
  const int smem = wg_smem(MG, NA);
  cudaError_t e;
  if (tf == TC_TF_SILU) {
    e = cudaFuncSetAttribute(tc_wgrad_kernel<MG, NA, TC_TF_SILU>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    tc_wgrad_kernel<MG, NA, TC_TF_SILU><<<grid, kThreads, smem, st>>>(tg, ta, n_stages, partial, cpart);
  } else {
    e = cudaFuncSetAttribute(tc_wgrad_kernel<MG, NA, TC_TF_NONE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    tc_wgrad_kernel<MG, NA, TC_TF_NONE><<<grid, kThreads, smem, st>>>(tg, ta, n_stages, partial, cpart);
  }
  AE_LAUNCH_CHECK();
  return 0;
}

}  // namespace

bool tc_wgrad_supported(int MG, int NA) {
  return (MG == 64 || MG == 128 || MG == 256) && (NA == 64 || NA == 128 || NA == 256);
}

size_t tc_wgrad_partial_floats(int MG, int NA) { return (size_t)tc_sm_count() * MG * (NA + 1); }

int tc_wgrad(const void* G, int64_t ldg, int MG, const void* ACT, int64_t lda, int NA, int tf, int64_t E, float* dW,
             int64_t lddw, int n_valid, int fold32, float* db, float* partial, size_t partial_floats, cudaStream_t st) {
  if (!tc_wgrad_supported(MG, NA)) return AETHER_E_UNSUPPORTED;
  const int64_t n_stages = (E + kStageRows - 1) / kStageRows;
  const int grid = wg_grid(n_stages);
  if (!partial || partial_floats < (size_t)grid * MG * (NA + 1)) return AETHER_E_WORKSPACE;
  float* cpart = db ? partial + (size_t)grid * MG * NA : nullptr;
  if (E > 0) {
    CUtensorMap tg, ta;
    AE_CHECK(tc_make_tmap_bf16(&tg, G, E, MG, ldg, kStageRows));
    AE_CHECK(tc_make_tmap_bf16(&ta, ACT, E, NA, lda, kStageRows));
    int rc = AETHER_E_UNSUPPORTED;
#define AE_WG_CASE(MM, NN) if (MG == MM && NA == NN) rc = launch_wg<MM, NN>(tg, ta, tf, n_stages, grid, partial, cpart, st);
    AE_WG_CASE(256, 256) AE_WG_CASE(256, 64) AE_WG_CASE(128, 128) AE_WG_CASE(128, 64) AE_WG_CASE(64, 64)
    AE_WG_CASE(256, 128) AE_WG_CASE(128, 256) AE_WG_CASE(64, 128) AE_WG_CASE(64, 256)
#undef AE_WG_CASE
    if (rc != 0) return rc;
  }
  const int tot = MG * n_valid;
  wgrad_reduce_kernel<<<(tot + 255) / 256, 256, 0, st>>>(partial, E > 0 ? grid : 0, MG, NA, n_valid, fold32, dW, lddw);
  AE_LAUNCH_CHECK();
  if (db) {
    colsum_reduce_kernel<<<(MG + 255) / 256, 256, 0, st>>>(cpart, E > 0 ? grid : 0, MG, db);
    AE_LAUNCH_CHECK();
  }
  return 0;
}

}  // namespace aether
