// Chained data-gradient GEMMs of the edge-MLP backward on a CTA pair: the step from layer l down to layer l-1 has no
// node-level dependency between its two GEMMs, so they run as ONE kernel (autograd of locs.py:232-235 / :206-212):
//
//     g2 = (g1_l W_e,l  (+ gyd[recv]))  *  silu'(pre2_{l-1})        dL/dpre2_{l-1}; gyd = gy_{l-1} / in-degree (the mean)
//     g1 = (g2 W_2,{l-1})               *  silu'(pre1_{l-1})        dL/dpre1_{l-1}
//
// g2 is stored (the weight gradient of W_2,{l-1} reads it) and, rounded to bf16, is written back into tensor memory over
// the accumulator's columns as the A operand of the second MMA -- the structure of the forward kernel (tc_chain.cu), same
// CTA pair, weight halves, TMEM plan, 16 warps with the setmaxnreg split.  What differs is where the epilogue operands
// come from: the silu' rows are two dense [E,H] streams from HBM, so they do NOT go through the LSU (row-per-thread loads
// touch 32 lines per instruction; that is what holds the separate kernels at 5.4 / 7.6 us per tile): the TMA producer
// loads them as 128x64 boxes into the SAME ring as the A operand -- per tile the ring carries A, Z1, Z2 in that order --
// and the epilogue warps read their rows out of the 128-byte-swizzled box (conflict-free 16-byte reads) and release the
// slot.  HBM passes per layer pair: 5 (g1, pre2, pre1 in; g2, g1 out) against 6 for the two kernels.
//
// Ring protocol.  Slot s = load index % S.  A-loads arm bar_raw[s], Z-loads arm bar_rawz[s] (TMA bytes) -- two barriers,
// because a parity wait cannot tell phase k from phase k+2: with one barrier per slot the relay warps (which only follow
// A-loads) and the epilogue warps (which only follow Z-loads) would each skip phases and alias.  A-loads are consumed by the MMA
// (bar_rdy relays "landed in both CTAs" to the leader, tcgen05.commit frees the slot on bar_free); Z-loads are consumed
// by this CTA's eight epilogue warps, each reading its 32x32 part of the box and arriving on bar_zfree[s] (count 8).  The
// producer, before reusing a slot, waits on the barrier of the slot's PREVIOUS use, whose kind it remembers in a bit mask;
// each barrier's parity is the count of its own uses of that slot.  All waits are bounded (trap, not hang).
#include <string.h>

#include "tc_gemm.cuh"

namespace aether {

namespace {

using namespace tc;

constexpr int kThreads = 512;                      // warpgroups: epilogue x2 | TMA, MMA, 2 idle | relay
constexpr int kEpiWarps = 8;
constexpr int kWarpTma = 8, kWarpMma = 9, kWarpRelay0 = 12;
constexpr int kRegsEpi = 192, kRegsAux = 64;
constexpr int kTileM = 128;
constexpr int kSlotBytes = kTileM * 128;
constexpr int kSmemMax = 232448;
constexpr int kStageBytes = kEpiWarps * 2 * 2048;
constexpr int kMiscBytes = 512;

__host__ __device__ constexpr int w_bytes(int H) { return H * (H / 2) * 2; }      // this CTA's half of one H x H image
__host__ __device__ constexpr int num_slots(int H) {
  int s = (kSmemMax - kMiscBytes - 2 * w_bytes(H) - kStageBytes) / kSlotBytes;
  return s > 8 ? 8 : s;
}
__host__ __device__ constexpr int smem_bytes(int H) { return 2 * w_bytes(H) + num_slots(H) * kSlotBytes + kStageBytes + kMiscBytes; }

// out = (v/2) + (v/2) w, w = 2 silu'(z) - 1 evaluated for a bf16 pair (the form of tc_gemm.cu's backward epilogue)
__device__ __forceinline__ uint32_t dsilu_mul_pack(float v0, float v1, uint32_t z) {
  const uint32_t w = bf2_dsilu_w(z);
  v0 *= 0.5f; v1 *= 0.5f;
  return pack_bf16(fmaf(v0, bf16_lo(w), v0), fmaf(v1, bf16_hi(w), v1));
}

template <int H, bool GATHER, bool MMA1>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
tc_chain_bwd_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmZ1,
                    const __grid_constant__ CUtensorMap tmZ2, const __grid_constant__ CUtensorMap tmO1,
                    const __grid_constant__ CUtensorMap tmO2, TcChainArgs g) {
  static_assert(H == 64 || H == 128 || H == 256, "unsupported shape");
  constexpr int KC = H / 64;                       // ring slots per operand tile
  constexpr int S = num_slots(H);
  constexpr int NP = H / 64;
  constexpr int NCH = H / 32;
  constexpr uint32_t IDESC = make_idesc(256, H, MAJOR_K, MAJOR_K);
  constexpr uint32_t TMEM_COLS = 2 * H;
  constexpr int WROW = (H / 2) * 128;
  constexpr int NOPS = MMA1 ? 3 : 2;               // operand tiles per edge tile in the ring: (A,) Z1, Z2
  static_assert(S >= 2, "pipeline too shallow");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  if (base & 1023u) __trap();
  const uint32_t sW1 = base;
  const uint32_t sW2 = sW1 + w_bytes(H);
  const uint32_t sA = sW2 + w_bytes(H);
  const uint32_t sStage = sA + S * kSlotBytes;
  const uint32_t sMisc = sStage + kStageBytes;
  const uint32_t bar_raw = sMisc;                  // [S] TMA landed, A operand                    (this CTA)
  const uint32_t bar_rdy = sMisc + 64;             // [S] A landed in both CTAs, 8 relay warps      (leader's is used)
  const uint32_t bar_free = sMisc + 128;           // [S] MMA 1 consumed the A slot                 (multicast to both)
  const uint32_t bar_zfree = sMisc + 192;          // [S] the 8 epilogue warps read the Z slot      (this CTA)
  const uint32_t bar_rawz = sMisc + 256;           // [S] TMA landed, silu' rows                    (this CTA)
  const uint32_t bar_a2 = sMisc + 320;             // [NP] piece of A2 in tensor memory, 16 warps   (leader's)
  const uint32_t bar_acc1f = sMisc + 352;
  const uint32_t bar_acc2f = sMisc + 360;
  const uint32_t bar_acc2e = sMisc + 368;          // accumulator 2 drained, 16 warps of the pair   (leader's)
  const uint32_t bar_w = sMisc + 376;
  const uint32_t tmem_slot = sMisc + 384;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int64_t n_pt = (g.M + 2 * kTileM - 1) / (2 * kTileM);
  const int64_t cid = blockIdx.x >> 1, ncl = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(bar_raw + 8 * s, 1);
      mbar_init(bar_rdy + 8 * s, 8);
      mbar_init(bar_free + 8 * s, 1);
      mbar_init(bar_zfree + 8 * s, kEpiWarps);
      mbar_init(bar_rawz + 8 * s, 1);
    }
    for (int h = 0; h < NP; ++h) mbar_init(bar_a2 + 8 * h, 2 * kEpiWarps);
    mbar_init(bar_acc1f, 1);
    mbar_init(bar_acc2f, 1);
    mbar_init(bar_acc2e, 2 * kEpiWarps);
    mbar_init(bar_w, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (warp == kWarpMma) tmem_alloc_2cta(tmem_slot, TMEM_COLS);
  if (warp == kWarpTma && lane == 0) {
    tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmZ1); tma_prefetch_desc(&tmZ2);
    tma_prefetch_desc(&tmO1); tma_prefetch_desc(&tmO2);
    mbar_arrive_expect_tx(bar_w, (MMA1 ? 2 : 1) * w_bytes(H));
    for (int c = 0; c < KC; ++c) {
      if (MMA1) bulk_load(sW1 + c * WROW, (const uint8_t*)g.wpack1 + (size_t)c * (H * 128) + (size_t)rank * WROW, WROW, bar_w);
      bulk_load(sW2 + c * WROW, (const uint8_t*)g.wpack2 + (size_t)c * (H * 128) + (size_t)rank * WROW, WROW, bar_w);
    }
    mbar_wait(bar_w, 0);
  }
  __syncwarp();
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  if (warp >= kEpiWarps) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" :: "n"(kRegsAux));

  if (warp == kWarpTma) {
    // ===== TMA producer: per tile A[0..KC), Z1[0..KC), Z2[0..KC) through one ring =====
    if (lane == 0) {
      uint32_t it = 0, nA = 0, nZ = 0, lastZ = 0;  // bit s: parity of the A- / Z-uses of slot s so far; kind of its last use
      for (int64_t pt = cid; pt < n_pt; pt += ncl) {
        const int row0 = (int)(pt * 2 * kTileM + rank * kTileM);
#pragma unroll 1
        for (int op = MMA1 ? 0 : 1; op < 3; ++op) {
          const CUtensorMap* tm = op == 0 ? &tmA : (op == 1 ? &tmZ1 : &tmZ2);
          for (int c = 0; c < KC; ++c, ++it) {
            const uint32_t s = it % S, bit = 1u << s;
            if (it >= (uint32_t)S) {               // the slot's previous use must have been consumed
              if (lastZ & bit) mbar_wait(bar_zfree + 8 * s, ((nZ >> s) & 1) ^ 1);
              else mbar_wait(bar_free + 8 * s, ((nA >> s) & 1) ^ 1);
            }
            const uint32_t land = (op == 0 ? bar_raw : bar_rawz) + 8 * s;
            mbar_arrive_expect_tx(land, kSlotBytes);
            tma_load_2d(sA + s * kSlotBytes, tm, c * 64, row0, land);
            if (op == 0) { nA ^= bit; lastZ &= ~bit; } else { nZ ^= bit; lastZ |= bit; }
          }
        }
      }
    }
  } else if (warp == kWarpMma) {
    // ===== MMA issuer (leader CTA) =====
    if (rank == 0 && lane == 0) {
      const uint32_t acc1 = tmem_base, acc2 = tmem_base + H;
      uint32_t lt = 0, nA = 0;
      for (int64_t pt = cid; pt < n_pt; pt += ncl, ++lt) {
        for (int c = 0; MMA1 && c < KC; ++c) {
          const uint32_t it = lt * NOPS * KC + c, s = it % S;
          mbar_wait_cluster(bar_rdy + 8 * s, (nA >> s) & 1);
          nA ^= 1u << s;
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t ad = make_sdesc(sA + s * kSlotBytes + k * 32, 16, 1024);
            const uint64_t bd = make_sdesc(sW1 + c * WROW + k * 32, 16, 1024);
            umma_bf16_2cta(acc1, ad, bd, IDESC, (c | k) != 0);
          }
          umma_commit_2cta(bar_free + 8 * s);
        }
        if (MMA1) umma_commit_2cta(bar_acc1f);
        mbar_wait_cluster(bar_acc2e, (lt & 1) ^ 1);
        tc_fence_after();
        for (int h = 0; h < NP; ++h) {
          mbar_wait_cluster(bar_a2 + 8 * h, lt & 1);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint32_t ks = 4 * h + k;
            const uint32_t a_t = acc1 + 32 * (ks >> 1) + 8 * (ks & 1);
            const uint64_t bd = make_sdesc(sW2 + h * WROW + k * 32, 16, 1024);
            umma_bf16_2cta_ts(acc2, a_t, bd, IDESC, (h | k) != 0);
          }
        }
        umma_commit_2cta(bar_acc2f);
      }
    }
  } else if (warp >= kWarpRelay0) {
    // ===== relay: "the A box of this CTA has landed" to the leader's barrier (the operand needs no transform) =====
    const uint32_t rdy0 = mapa_rank(bar_rdy, 0);
    uint32_t lt = 0, nA = 0;
    for (int64_t pt = cid; pt < n_pt; pt += ncl, ++lt) {
      for (int c = 0; MMA1 && c < KC; ++c) {
        const uint32_t it = lt * NOPS * KC + c, s = it % S;
        mbar_wait(bar_raw + 8 * s, (nA >> s) & 1);
        nA ^= 1u << s;
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(rdy0 + 8 * s);
      }
    }
  } else if (warp < kEpiWarps) {
    // ===== epilogue warps =====
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" :: "n"(kRegsEpi));
    constexpr int CPW = NCH / 2;                   // chunks per warp and epilogue phase
    constexpr int D = CPW >= 2 ? 2 : 1;            // gyd[recv] rows are requested D chunks ahead
    constexpr bool TOP = (D % CPW) != 0;
    const int qr = warp & 3, hf = warp >> 2;
    const uint32_t sbuf = sStage + warp * 4096;
    const uint32_t a2bar0 = mapa_rank(bar_a2, 0), acc2e0 = mapa_rank(bar_acc2e, 0);
    const uint32_t lane_base = tmem_base + ((uint32_t)(qr * 32) << 16);
    const int trow = qr * 32 + lane;               // this thread's row of the CTA's 128-row tile
    const uint32_t zrow = (uint32_t)((trow >> 3) * 1024 + (trow & 7) * 128);
    uint32_t lt = 0, nbox = 0, nZ = 0;             // nZ bit s: parity of the Z boxes this warp has taken from slot s
    uint4 gr[CPW][4];
    int ib = 0, ib_n = 0;
    auto request = [&](uint4 (&dr)[4], int ib_, int c) {
      const uint4* pb = reinterpret_cast<const uint4*>(g.gat16 + ((int64_t)ib_ * g.ldg + c));
#pragma unroll
      for (int q = 0; q < 4; ++q) dr[q] = __ldg(pb + q);
    };
    // this thread's 32 columns of a landed Z box (64 columns wide): 16-byte pieces 4*hf .. 4*hf+3 of its row
    auto read_z = [&](uint32_t slot, uint4 (&z)[4]) {
      const uint32_t a = sA + slot * kSlotBytes + zrow;
#pragma unroll
      for (int q = 0; q < 4; ++q)
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(z[q].x), "=r"(z[q].y), "=r"(z[q].z), "=r"(z[q].w)
                     : "r"(a + (uint32_t)((((4 * hf + q) ^ (trow & 7))) << 4)));
    };
    if (GATHER && cid < n_pt) {
      const int64_t r0 = cid * 2 * kTileM + rank * kTileM + trow;
      if (r0 < g.M) ib_n = g.idx_b[r0];
#pragma unroll
      for (int k = 0; k < D; ++k) request(gr[k], ib_n, (hf + 2 * k) * 32);
    }
    for (int64_t pt = cid; pt < n_pt; pt += ncl, ++lt) {
      const int64_t row0 = pt * 2 * kTileM + rank * kTileM + qr * 32;
      const int64_t ptn = pt + ncl;
      const int64_t row_n = row0 + lane + ncl * 2 * kTileM;
      ib = ib_n;
      ib_n = 0;
      if (GATHER && ptn < n_pt && row_n < g.M) ib_n = g.idx_b[row_n];
      auto prefetch = [&](int kk) {
        const int kn = kk + D;
        if (kn < CPW) request(gr[kn % CPW], ib, (hf + 2 * kn) * 32);
        else if (ptn < n_pt) request(gr[kn % CPW], ib_n, (hf + 2 * (kn - CPW)) * 32);
      };

      // ---- epilogue 1: g2 = (acc1 + gyd[recv]) * silu'(Z1) -> HBM and, as bf16, back into tensor memory (A2) ----
      // (without the first MMA -- the top layer, g2 = gyd[recv] * silu'(Z1) -- accumulator 1's columns are free once MMA 2 of
      // the previous tile has completed, which epilogue 2 of that tile already waited for)
      if (MMA1) mbar_wait(bar_acc1f, lt & 1);
      tc_fence_after();
#pragma unroll
      for (int kk = 0; kk < CPW; ++kk) {
        const int j = hf + 2 * kk;
        const int c0 = j * 32;
        if (GATHER && TOP) prefetch(kk);
        uint32_t r[32];
        if (MMA1) {
          tmem_ld32(lane_base + c0, r);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = 0u;
        }
        const uint32_t itz = lt * NOPS * KC + (NOPS - 2) * KC + kk, zs = itz % S;
        mbar_wait(bar_rawz + 8 * zs, (nZ >> zs) & 1);
        nZ ^= 1u << zs;
        uint4 z[4];
        read_z(zs, z);
        const uint32_t buf = sbuf + (nbox & 1) * 2048;
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        __syncwarp();
        tmem_ld_wait();
        uint32_t o[16];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t zz[4] = {z[q].x, z[q].y, z[q].z, z[q].w};
          const uint32_t gg[4] = {gr[kk][q].x, gr[kk][q].y, gr[kk][q].z, gr[kk][q].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float v0 = __uint_as_float(r[8 * q + 2 * e]), v1 = __uint_as_float(r[8 * q + 2 * e + 1]);
            if (GATHER) { v0 += bf16_lo(gg[e]); v1 += bf16_hi(gg[e]); }
            o[4 * q + e] = dsilu_mul_pack(v0, v1, zz[e]);
          }
        }
        if (GATHER && !TOP) prefetch(kk);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t addr = buf + lane * 64 + ((q ^ ((lane >> 1) & 3)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(o[4 * q]), "r"(o[4 * q + 1]),
                       "r"(o[4 * q + 2]), "r"(o[4 * q + 3]) : "memory");
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(bar_zfree + 8 * zs);         // every lane's Z values have been consumed above
          if (row0 < g.M)
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                         :: "l"((uint64_t)&tmO1), "r"(c0), "r"((int)row0), "r"(buf) : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        ++nbox;
        tmem_st16(lane_base + c0, o);              // A2 = bf16(g2): no activation between the two GEMMs
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(a2bar0 + 8 * (j >> 1));
      }

      // ---- epilogue 2: g1 = acc2 * silu'(Z2) -> HBM ----
      mbar_wait(bar_acc2f, lt & 1);
      tc_fence_after();
      {
        uint32_t r[32];
        tmem_ld32(lane_base + H + hf * 32, r);
#pragma unroll
        for (int kk = 0; kk < CPW; ++kk, ++nbox) {
          const int c0 = (hf + 2 * kk) * 32;
          const uint32_t buf = sbuf + (nbox & 1) * 2048;
          const uint32_t itz = lt * NOPS * KC + (NOPS - 1) * KC + kk, zs = itz % S;
          mbar_wait(bar_rawz + 8 * zs, (nZ >> zs) & 1);
          nZ ^= 1u << zs;
          uint4 z[4];
          read_z(zs, z);
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          tmem_ld_wait();
          uint32_t o[16];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t zz[4] = {z[q].x, z[q].y, z[q].z, z[q].w};
#pragma unroll
            for (int e = 0; e < 4; ++e)
              o[4 * q + e] = dsilu_mul_pack(__uint_as_float(r[8 * q + 2 * e]), __uint_as_float(r[8 * q + 2 * e + 1]), zz[e]);
          }
          if (kk + 1 < CPW) {
            tmem_ld32(lane_base + H + c0 + 64, r);
          } else {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(acc2e0);
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t addr = buf + lane * 64 + ((q ^ ((lane >> 1) & 3)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(o[4 * q]), "r"(o[4 * q + 1]),
                         "r"(o[4 * q + 2]), "r"(o[4 * q + 3]) : "memory");
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(bar_zfree + 8 * zs);
            if (row0 < g.M)
              asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                           :: "l"((uint64_t)&tmO2), "r"(c0), "r"((int)row0), "r"(buf) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        }
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");

  }
  __syncwarp();
  tc_fence_before();
  cluster_sync_all();
  if (warp == kWarpMma) {
    tc_fence_after();
    tmem_dealloc_2cta(tmem_base, TMEM_COLS);
  }

}

template <int H, bool GATHER, bool MMA1>
int launch_bwd(const TcChainArgs& g, const CUtensorMap& ta, const CUtensorMap& z1, const CUtensorMap& z2, const CUtensorMap& o1,
               const CUtensorMap& o2, cudaStream_t st) {
  constexpr int smem = smem_bytes(H);
  static_assert(smem <= kSmemMax, "shared memory budget");
  auto kern = tc_chain_bwd_kernel<H, GATHER, MMA1>;
  static int max_clusters = 0;
  if (max_clusters == 0) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(2 * (unsigned)tc_sm_count());
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at;
    at.id = cudaLaunchAttributeClusterDimension;
    at.val.clusterDim.x = 2; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
    cfg.attrs = &at; cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) { (void)cudaGetLastError(); n = tc_sm_count() / 2; }
    max_clusters = n;
  }
  const int64_t n_pt = (g.M + 2 * kTileM - 1) / (2 * kTileM);
  const unsigned ncl = (unsigned)(n_pt < max_clusters ? n_pt : max_clusters);
  kern<<<2 * ncl, kThreads, smem, st>>>(ta, z1, z2, o1, o2, g);
  AE_LAUNCH_CHECK();
  return 0;
}

}  // namespace

bool tc_chain_bwd_supported(int H) { return H == 64 || H == 128 || H == 256; }

// g.wpack1 / g.wpack2: TRANSPOSED images (tc_pack_weight transpose = 1) of W_e,l and W_2,{l-1}; g.gat16 + g.idx_b: the
// receiver-indexed node term (optional); z1 / z2: saved pre-activations; g.out1 = g2, g.out2 = g1 (both required).
// A == nullptr: no first GEMM (top layer): out1 = gat16[idx_b[m]] * silu'(z1), g.wpack1 is not read.
int tc_chain_bwd(const TcChainArgs& g, const void* A, int64_t lda, const void* z1, int64_t ldz1, const void* z2, int64_t ldz2,
                 int H, cudaStream_t st) {
  if (g.M <= 0) return 0;
  if (!tc_chain_bwd_supported(H)) return AETHER_E_UNSUPPORTED;
  if (!g.wpack2 || !g.out1 || !g.out2 || !z1 || !z2 || g.bias1 || g.bias2 || g.out2_f32) return AETHER_E_BADARG;
  if (A ? !g.wpack1 : !g.gat16) return AETHER_E_BADARG;
  if (g.gat16 && (!g.idx_b || ((uintptr_t)g.gat16 & 15) || (g.ldg & 7))) return AETHER_E_BADARG;
  CUtensorMap ta, tz1, tz2, t1, t2;
  AE_CHECK(tc_make_tmap_bf16(&ta, A ? A : z1, g.M, H, A ? lda : ldz1, kTileM));
  AE_CHECK(tc_make_tmap_bf16(&tz1, z1, g.M, H, ldz1, kTileM));
  AE_CHECK(tc_make_tmap_bf16(&tz2, z2, g.M, H, ldz2, kTileM));
  AE_CHECK(tc_make_tmap_bf16(&t1, g.out1, g.M, H, g.ldo1, 32, 32));
  AE_CHECK(tc_make_tmap_bf16(&t2, g.out2, g.M, H, g.ldo2, 32, 32));
  const bool gather = g.gat16 != nullptr;
#define AE_CB_CASE(HH)                                                                                                  \
  if (H == HH) {                                                                                                        \
    if (!A) return launch_bwd<HH, true, false>(g, ta, tz1, tz2, t1, t2, st);                                            \
    return gather ? launch_bwd<HH, true, true>(g, ta, tz1, tz2, t1, t2, st) : launch_bwd<HH, false, true>(g, ta, tz1, tz2, t1, t2, st); \
  }
  AE_CB_CASE(256) AE_CB_CASE(128) AE_CB_CASE(64)
#undef AE_CB_CASE
  return AETHER_E_UNSUPPORTED;
}

}  // namespace aether
