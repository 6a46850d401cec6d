// Chained edge MLP of one message-passing layer on a CTA pair (nn/state2state/locs/locs.py:206-212,
// 232-235):
//
//     pre1 = tf(A) We^T + b1 (+ P_s[send] + P_r[recv])          A = stored pre2 of the previous layer (or the features)
//     pre2 = silu(pre1) W2^T + b2
//
// ONE persistent kernel, the intermediate never leaves the SM: accumulator 1 is read from tensor memory, finished in
// registers, passed through SiLU, packed to bf16 and written BACK into tensor memory over the columns it came from,
// where it is the A operand of the second tcgen05.mma (A from TMEM).  pre1 goes to HBM only when the backward needs
// it (SAVE); pre2 always does (it is the layer's edge state).  HBM traffic per edge: 2*K1 bytes in, 2*H (+ 2*H) out,
// against 2*K1 + 6*H for the two separate GEMMs.
//
// Both weight images do not fit one SM at H = 256 (2 x 128 KB), so the kernel runs on CTA PAIRS (cluster of two,
// tcgen05 cta_group::2): an MMA covers 256 edges (128 per CTA), each CTA holds HALF the rows of W_e and of W_2
// (64 KB + 64 KB at H = 256) and the tensor cores of both SMs read both halves.  Per CTA:
//
//   warp 8      TMA producer : its halves of the two weight images once, then 128x64 bf16 boxes of A (ring)
//   warps 12-15 transform    : in-place SiLU on the landed box, then "ready" to the LEADER's barrier (remote arrive)
//   warp 9      MMA issuer   : CTA 0 only; MMA 1 (A, B from shared memory) into accumulator 1, MMA 2 (A from
//                              tensor memory, B from shared memory) into accumulator 2; completion is multicast
//                              to the barriers of both CTAs (tcgen05.commit ... multicast::cluster)
//   warps 0-7   epilogue     : two warps per TMEM lane quarter on alternate 32-column chunks
//        epilogue 1: tcgen05.ld, + b1 + gathered node terms, bf16 pre1 -> TMA store (SAVE), SiLU, bf16 pair per
//                    32-bit column -> tcgen05.st over columns [32j, 32j+16) of the chunk, "piece ready" to the
//                    leader after each chunk (MMA 2 starts on the first 64 inputs while the rest of the row is
//                    still in the epilogue)
//        epilogue 2: tcgen05.ld (the next chunk's load in flight under this chunk's stores), + b2, bf16 pre2 -> TMA store
//
// 16 warps = four complete warpgroups, so the register file is re-split after the set-up (setmaxnreg): the two
// epilogue warpgroups take 192 registers per thread, the producer / MMA / transform warpgroups keep 64.  The larger
// budget holds one register slot per chunk position for the gathered rows P_s[send], P_r[recv] (64 bytes each per
// row and chunk), requested TWO chunks ahead -- also across tiles -- with static slot indices (no register copies
// that would wait on the loads).  profiles/r02_ncu_chain_v1.md: with one chunk ahead the epilogue warps sat on
// exactly these loads (gather variant 18.4k cycles per tile against 12.3k without).
//
// Tensor memory: accumulator 1 at columns [0, H), accumulator 2 at [H, 2H) -- all 512 columns at H = 256, so there
// is no second accumulator pair; MMA 1 of the next tile runs while epilogue 2 of this one drains.
// Numerics are those of the two separate kernels (same operands, same rounding points, same K order): the
// outputs are bit-identical to tc_gemm + tc_gemm, which tests/test_chain_gpu.py checks.
#include <string.h>

#include "tc_gemm.cuh"

namespace aether {

namespace {

using namespace tc;

constexpr int kThreads = 512;                      // warpgroups: epilogue x2 | TMA, MMA, 2 idle | transform
constexpr int kEpiWarps = 8;
constexpr int kWarpTma = 8, kWarpMma = 9, kWarpTf0 = 12;
constexpr int kRegsEpi = 192, kRegsAux = 64;       // 8 x 32 x 192 + 8 x 32 x 64 = 64 Ki registers
constexpr int kTileM = 128;                        // rows per CTA; the pair owns 256
constexpr int kSlotBytes = kTileM * 128;           // 128 rows x 64 bf16
constexpr int kSmemMax = 232448;                   // 227 KB
constexpr int kStageBytes = kEpiWarps * 2 * 2048;  // per epilogue warp: two 32x32 bf16 boxes

__host__ __device__ constexpr int misc_bytes(int H) { return 512 + 2 * H * 4; }      // barriers | b1 | b2
__host__ __device__ constexpr int w1_bytes(int K1, int H) { return K1 * (H / 2) * 2; }  // this CTA's half of the rows
__host__ __device__ constexpr int w2_bytes(int H) { return H * (H / 2) * 2; }
__host__ __device__ constexpr int num_slots(int K1, int H) {
  int s = (kSmemMax - misc_bytes(H) - w1_bytes(K1, H) - w2_bytes(H) - kStageBytes) / kSlotBytes;
  return s > 8 ? 8 : s;
}
__host__ __device__ constexpr int smem_bytes(int K1, int H) {
  return w1_bytes(K1, H) + w2_bytes(H) + num_slots(K1, H) * kSlotBytes + kStageBytes + misc_bytes(H);
}

template <int K1, int H, int TF1, bool SAVE, bool GATHER, bool OUT32, bool BOX>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
tc_chain_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmO1,
                const __grid_constant__ CUtensorMap tmO2, const __grid_constant__ CUtensorMap tmG, TcChainArgs g) {
  static_assert(!BOX || GATHER, "the box feed is a form of the gather variant");
  static_assert(K1 % 64 == 0 && (H == 64 || H == 128 || H == 256), "unsupported shape");
  constexpr int KC1 = K1 / 64;                     // 64-wide K chunks of the first GEMM = ring slots per tile
  constexpr int S = num_slots(K1, H);
  constexpr int NP = H / 64;                       // 64-input pieces of the second GEMM
  constexpr int NCH = H / 32;                      // 32-column epilogue chunks
  constexpr uint32_t IDESC = make_idesc(256, H, MAJOR_K, MAJOR_K);
  constexpr uint32_t TMEM_COLS = 2 * H;            // 128 / 256 / 512
  constexpr int WROW = (H / 2) * 128;              // bytes of one 64-wide K chunk of this CTA's weight half
  // BOX: the sender term P_s[send] of a tile whose senders span fewer than 128 node rows (batched small graphs: a tile
  // lies inside one graph) is ONE 128-row box of the node-term table per 64 columns, loaded by TMA into the same ring
  // after the A boxes; the epilogue reads row (send - base) of the box.  Tiles that do not fit (graph boundaries) fall
  // back to direct loads.  Chosen per launch from the graph's statistics (TcChainArgs::box_gather).
  constexpr int NG = BOX ? H / 64 : 0;             // sender-term boxes per tile
  constexpr int LPT = KC1 + NG;                    // ring loads per tile
  static_assert(S >= 2, "pipeline too shallow");

  // The dynamic shared memory of a kernel without static shared memory starts 1024-byte aligned (the swizzle atom of
  // the operand tiles); no slack is reserved for rounding up, which is what makes room for the fourth ring slot at
  // H = 256 (a whole tile of A in flight: with three, MMA 1 of the next tile waited on its last chunk's TMA + SiLU).
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  if (base & 1023u) __trap();
  const uint32_t sW1 = base;
  const uint32_t sW2 = sW1 + w1_bytes(K1, H);
  const uint32_t sA = sW2 + w2_bytes(H);
  const uint32_t sStage = sA + S * kSlotBytes;
  const uint32_t sMisc = sStage + kStageBytes;
  const uint32_t bar_raw = sMisc;                  // [S] TMA landed                       (this CTA)
  const uint32_t bar_rdy = sMisc + 64;             // [S] transform done, 8 warps of the pair    (leader's is used)
  const uint32_t bar_free = sMisc + 128;           // [S] MMA 1 consumed the slot          (multicast to both)
  const uint32_t bar_a2 = sMisc + 192;             // [NP] piece of A2 in tensor memory, 16 warps of the pair (leader's)
  const uint32_t bar_acc1f = sMisc + 224;          // accumulator 1 complete               (multicast to both)
  const uint32_t bar_acc2f = sMisc + 232;          // accumulator 2 complete               (multicast to both)
  const uint32_t bar_acc2e = sMisc + 240;          // accumulator 2 drained, 16 warps of the pair        (leader's)
  const uint32_t bar_w = sMisc + 248;              // weights landed                       (this CTA)
  const uint32_t tmem_slot = sMisc + 256;
  const uint32_t bar_zfree = sMisc + 320;          // [S] the 8 epilogue warps read the sender-term box   (this CTA)
  const uint32_t bar_rawz = sMisc + 384;           // [S] sender-term box landed (or: tile falls back)    (this CTA)
  const uint32_t sBase = sMisc + 448;              // [S] int32: first node row of the box in the slot, -1 = fall back
  uint8_t* misc_ptr = smem_raw + (sMisc - smem_u32(smem_raw));
  float* sb1 = reinterpret_cast<float*>(misc_ptr + 512);
  float* sb2 = sb1 + H;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int64_t n_pt = (g.M + 2 * kTileM - 1) / (2 * kTileM);     // pair tiles of 256 rows
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
  if (threadIdx.x < H) {
    sb1[threadIdx.x] = g.bias1 ? g.bias1[threadIdx.x] : 0.0f;
    sb2[threadIdx.x] = g.bias2 ? g.bias2[threadIdx.x] : 0.0f;
  }
  __syncthreads();
  if (warp == kWarpMma) tmem_alloc_2cta(tmem_slot, TMEM_COLS);
  if (warp == kWarpTma && lane == 0) {
    tma_prefetch_desc(&tmA);
    if (BOX) tma_prefetch_desc(&tmG);
    if (!OUT32) tma_prefetch_desc(&tmO2);
    if (SAVE) tma_prefetch_desc(&tmO1);
    // this CTA's half of the rows of every 64-wide K chunk of both images
    mbar_arrive_expect_tx(bar_w, w1_bytes(K1, H) + w2_bytes(H));
    for (int c = 0; c < KC1; ++c)
      bulk_load(sW1 + c * WROW, (const uint8_t*)g.wpack1 + (size_t)c * (H * 128) + (size_t)rank * WROW, WROW, bar_w);
    for (int c = 0; c < NP; ++c)
      bulk_load(sW2 + c * WROW, (const uint8_t*)g.wpack2 + (size_t)c * (H * 128) + (size_t)rank * WROW, WROW, bar_w);
    mbar_wait(bar_w, 0);
  }
  __syncwarp();
  tc_fence_before();
  cluster_sync_all();            // barriers initialised, weights resident and TMEM allocated in BOTH CTAs
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  // one instruction for both auxiliary warpgroups; the epilogue warpgroups raise theirs inside their own branch
  if (warp >= kEpiWarps) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" :: "n"(kRegsAux));

  if (warp == kWarpTma) {
    // ===== TMA producer (both CTAs, own rows): per tile the A boxes, then (BOX) the sender-term boxes.  Ring protocol as
    // in tc_chain_bwd.cu: before reusing a slot the producer waits on the barrier of the slot's previous use (MMA commit
    // for A boxes, the eight epilogue warps for sender-term boxes); separate landing barriers per kind. =====
    uint32_t it = 0, nA = 0, nZ = 0, lastZ = 0;
    for (int64_t pt = cid; pt < n_pt; pt += ncl) {
      const int row0 = (int)(pt * 2 * kTileM + rank * kTileM);
      int base = -1;
      if (BOX) {
        // senders of this CTA's 128 rows: 4 per lane; the box fits if they span fewer than 128 node rows
        int lo = 0x7fffffff, hi = -1;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int64_t r = (int64_t)row0 + 4 * lane + q;
          if (r < g.M) { const int v = g.idx_a[r]; lo = v < lo ? v : lo; hi = v > hi ? v : hi; }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
          const int l2 = __shfl_xor_sync(0xffffffffu, lo, o), h2 = __shfl_xor_sync(0xffffffffu, hi, o);
          lo = l2 < lo ? l2 : lo; hi = h2 > hi ? h2 : hi;
        }
        if (hi >= 0 && hi - lo < kTileM) base = lo;
      }
      if (lane == 0) {
        for (int c = 0; c < LPT; ++c, ++it) {
          const uint32_t s = it % S, bit = 1u << s;
          const bool isA = c < KC1;
          if (it >= (uint32_t)S) {
            if (lastZ & bit) mbar_wait(bar_zfree + 8 * s, ((nZ >> s) & 1) ^ 1);
            else mbar_wait(bar_free + 8 * s, ((nA >> s) & 1) ^ 1);
          }
          if (isA) {
            mbar_arrive_expect_tx(bar_raw + 8 * s, kSlotBytes);
            tma_load_2d(sA + s * kSlotBytes, &tmA, c * 64, row0, bar_raw + 8 * s);
            nA ^= bit; lastZ &= ~bit;
          } else {
            asm volatile("st.shared.b32 [%0], %1;" :: "r"(sBase + 4 * s), "r"(base) : "memory");
            if (base >= 0) {
              mbar_arrive_expect_tx(bar_rawz + 8 * s, kSlotBytes);
              tma_load_2d(sA + s * kSlotBytes, &tmG, (c - KC1) * 64, base, bar_rawz + 8 * s);
            } else {
              mbar_arrive(bar_rawz + 8 * s);       // nothing to land: the epilogue loads this tile's rows itself
            }
            nZ ^= bit; lastZ |= bit;
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
        // accumulator 1 is free: MMA 2 of the previous tile (which read A2 out of its columns) was issued before
        // this point and the tensor pipe executes in issue order
        for (int c = 0; c < KC1; ++c) {
          const uint32_t s = (lt * LPT + c) % S;
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
        umma_commit_2cta(bar_acc1f);
        mbar_wait_cluster(bar_acc2e, (lt & 1) ^ 1);
        tc_fence_after();
        for (int h = 0; h < NP; ++h) {
          mbar_wait_cluster(bar_a2 + 8 * h, lt & 1);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // K step ks = 4h + k: 16 inputs = 8 packed columns at the start of 32-column chunk ks/2, second half if odd
            const uint32_t ks = 4 * h + k;
            const uint32_t a_t = acc1 + 32 * (ks >> 1) + 8 * (ks & 1);
            const uint64_t bd = make_sdesc(sW2 + h * WROW + k * 32, 16, 1024);
            umma_bf16_2cta_ts(acc2, a_t, bd, IDESC, (h | k) != 0);
          }
        }
        umma_commit_2cta(bar_acc2f);
      }
    }
  } else if (warp >= kWarpTf0) {
    // ===== transform: in-place SiLU on the landed A box, then "ready" to the leader =====
    const int tt = threadIdx.x - kWarpTf0 * 32;    // 0..127
    const uint32_t rdy0 = mapa_rank(bar_rdy, 0);
    uint32_t lt = 0, nA = 0;
    for (int64_t pt = cid; pt < n_pt; pt += ncl, ++lt) {
      for (int c = 0; c < KC1; ++c) {
        const uint32_t s = (lt * LPT + c) % S;
        mbar_wait(bar_raw + 8 * s, (nA >> s) & 1);
        nA ^= 1u << s;
        if (TF1 == TC_TF_SILU) {
          const uint32_t a0 = sA + s * kSlotBytes + tt * 16;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            uint32_t u[4];
            const uint32_t addr = a0 + i * 2048;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]) : "r"(addr));
#pragma unroll
            for (int q = 0; q < 4; ++q) u[q] = pack_bf16(silu_fast(bf16_lo(u[q])), silu_fast(bf16_hi(u[q])));
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[3]) : "memory");
          }
          fence_proxy_async_smem();
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(rdy0 + 8 * s);
      }
    }
  } else if (warp < kEpiWarps) {
    // ===== epilogue warps =====
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" :: "n"(kRegsEpi));
    constexpr int CPW = NCH / 2;                   // chunks per warp and epilogue phase: 4 / 2 / 1
    constexpr int D = CPW >= 2 ? 2 : 1;            // the gathered rows are requested D chunks ahead
    constexpr bool TOP = (D % CPW) != 0;           // the target slot is not the one being consumed: request first
    const int qr = warp & 3, hf = warp >> 2;
    const uint32_t sbuf = sStage + warp * 4096;
    const uint32_t a2bar0 = mapa_rank(bar_a2, 0), acc2e0 = mapa_rank(bar_acc2e, 0);
    const uint32_t lane_base = tmem_base + ((uint32_t)(qr * 32) << 16);
    const bool has_b1 = g.bias1 != nullptr;
    uint32_t lt = 0, nbox = 0, nZ = 0;             // nZ bit s: parity of the sender-term boxes this warp took from slot s
    // gathered node terms: slot kk holds the 64-byte pieces of P_s[send] and P_r[recv] for chunk position kk
    uint4 gs[CPW][4], gr[CPW][4];
    int ia = 0, ib = 0, ia_n = 0, ib_n = 0;
    auto request = [&](uint4 (&ds)[4], uint4 (&dr)[4], int ia_, int ib_, int c) {
      // 128-bit loads: the 256-bit form (ld.global.nc.v8.b32) measured the same on contiguous senders and 16 % slower
      // on random ones (profiles/r02_chain_bench_v3.txt).  Feeding P_s[send] through the A ring with the TMA row gather
      // (cp.async.bulk.tensor tile::gather4, 32 instructions per 128 x 64 box; it works -- tools/gather4_probe.py -- and
      // the results were bit-identical) measured SLOWER, 9.8 vs 8.6 us per tile whatever the index pattern
      // (profiles/r02_chain_bench_gather4.txt): ~58 cycles per gather4, and the boxes can only be requested once MMA 1
      // has freed the slots.  The dense silu' operands of the backward kernel do go through the ring (tc_chain_bwd.cu).
      const uint4* pa = reinterpret_cast<const uint4*>(g.gat16 + (size_t)ia_ * g.ldg + c);
      const uint4* pb = reinterpret_cast<const uint4*>(g.gat16 + ((int64_t)ib_ * g.ldg + g.goff_b + c));
#pragma unroll
      for (int q = 0; q < 4; ++q) { if (!BOX) ds[q] = __ldg(pa + q); dr[q] = __ldg(pb + q); }
    };
    if (GATHER && cid < n_pt) {
      const int64_t r0 = cid * 2 * kTileM + rank * kTileM + qr * 32 + lane;
      if (r0 < g.M) { ia_n = g.idx_a[r0]; ib_n = g.idx_b[r0]; }
#pragma unroll
      for (int k = 0; k < D; ++k) request(gs[k], gr[k], ia_n, ib_n, (hf + 2 * k) * 32);
    }
    for (int64_t pt = cid; pt < n_pt; pt += ncl, ++lt) {
      const int64_t row0 = pt * 2 * kTileM + rank * kTileM + qr * 32;
      const int64_t ptn = pt + ncl;
      const int64_t row_n = row0 + lane + ncl * 2 * kTileM;
      ia = ia_n; ib = ib_n;
      ia_n = 0; ib_n = 0;
      if (GATHER && ptn < n_pt && row_n < g.M) { ia_n = g.idx_a[row_n]; ib_n = g.idx_b[row_n]; }
      auto prefetch = [&](int kk) {                // rows of the chunk D positions after (this tile, kk)
        const int kn = kk + D;
        if (kn < CPW) request(gs[kn % CPW], gr[kn % CPW], ia, ib, (hf + 2 * kn) * 32);
        else if (ptn < n_pt) request(gs[kn % CPW], gr[kn % CPW], ia_n, ib_n, (hf + 2 * (kn - CPW)) * 32);
      };

      // ---- epilogue 1: accumulator 1 -> pre1 (HBM, SAVE) and A2 = bf16(silu(pre1)) (tensor memory) ----
      mbar_wait(bar_acc1f, lt & 1);
      tc_fence_after();
#pragma unroll
      for (int kk = 0; kk < CPW; ++kk) {
        const int j = hf + 2 * kk;
        const int c0 = j * 32;
        if (GATHER && TOP) prefetch(kk);
        uint32_t r[32];
        tmem_ld32(lane_base + c0, r);
        f32x2 bv[16];
        if (!GATHER) {
          // bias of the chunk into registers while the tensor-memory load is in flight
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 t4 = *reinterpret_cast<const float4*>(sb1 + c0 + 4 * q);
            bv[2 * q] = f2_make(t4.x, t4.y); bv[2 * q + 1] = f2_make(t4.z, t4.w);
          }
        }
        const uint32_t buf = sbuf + (nbox & 1) * 2048;
        if (SAVE) {
          // the TMA store that read this staging box two boxes ago must have finished reading it
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
        }
        tmem_ld_wait();
        f32x2 v[16];
        if (!GATHER) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = f2_add(f2_make_u(r[2 * i], r[2 * i + 1]), bv[i]);
        } else {
          if (has_b1) {      // the step folds b1 into P_r (node GEMM bias), so the gather variant normally has none
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float4 t4 = *reinterpret_cast<const float4*>(sb1 + c0 + 4 * q);
              v[2 * q] = f2_add(f2_make_u(r[4 * q], r[4 * q + 1]), f2_make(t4.x, t4.y));
              v[2 * q + 1] = f2_add(f2_make_u(r[4 * q + 2], r[4 * q + 3]), f2_make(t4.z, t4.w));
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = f2_make_u(r[2 * i], r[2 * i + 1]);
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t bb[4] = {gr[kk][q].x, gr[kk][q].y, gr[kk][q].z, gr[kk][q].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) v[4 * q + e] = f2_add(v[4 * q + e], f2_from_bf16x2(bb[e]));
          }
          if (BOX) {
            // sender term out of the box in ring slot zs (this thread's row is node `ia`, box row ia - base), or -- tiles
            // whose senders span 128 rows or more -- straight from the table
            const uint32_t zs = (lt * LPT + KC1 + kk) % S;
            mbar_wait(bar_rawz + 8 * zs, (nZ >> zs) & 1);
            nZ ^= 1u << zs;
            int base;
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(base) : "r"(sBase + 4 * zs) : "memory");
            uint4 a4[4];
            if (base >= 0) {
              int lr = ia - base;
              lr = lr < 0 ? 0 : (lr > kTileM - 1 ? kTileM - 1 : lr);          // rows past M carry no valid index
              const uint32_t za = sA + zs * kSlotBytes + (uint32_t)((lr >> 3) * 1024 + (lr & 7) * 128);
#pragma unroll
              for (int q = 0; q < 4; ++q)
                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a4[q].x), "=r"(a4[q].y), "=r"(a4[q].z), "=r"(a4[q].w)
                             : "r"(za + (uint32_t)(((4 * hf + q) ^ (lr & 7)) << 4)));
            } else {
              const uint4* pa = reinterpret_cast<const uint4*>(g.gat16 + (size_t)ia * g.ldg + c0);
#pragma unroll
              for (int q = 0; q < 4; ++q) a4[q] = __ldg(pa + q);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint32_t aa[4] = {a4[q].x, a4[q].y, a4[q].z, a4[q].w};
#pragma unroll
              for (int e = 0; e < 4; ++e) v[4 * q + e] = f2_add(v[4 * q + e], f2_from_bf16x2(aa[e]));
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_zfree + 8 * zs);      // the adds above consumed every lane's loads
          } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint32_t aa[4] = {gs[kk][q].x, gs[kk][q].y, gs[kk][q].z, gs[kk][q].w};
#pragma unroll
              for (int e = 0; e < 4; ++e) v[4 * q + e] = f2_add(v[4 * q + e], f2_from_bf16x2(aa[e]));
            }
          }
          if (!TOP) prefetch(kk);
        }
        uint32_t o[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) o[i] = f2_pack_bf16(v[i]);
        if (SAVE) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t addr = buf + lane * 64 + ((q ^ ((lane >> 1) & 3)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(o[4 * q]), "r"(o[4 * q + 1]),
                         "r"(o[4 * q + 2]), "r"(o[4 * q + 3]) : "memory");
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0 && row0 < g.M) {
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                         :: "l"((uint64_t)&tmO1), "r"(c0), "r"((int)row0), "r"(buf) : "memory");
          }
          if (lane == 0) asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          ++nbox;
        }
        // the second GEMM consumes what the backward will recompute from the STORED pre1: silu of the bf16 value
#pragma unroll
        for (int i = 0; i < 16; ++i) o[i] = f2_pack_bf16(f2_silu(f2_from_bf16x2(o[i])));
        tmem_st16(lane_base + c0, o);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(a2bar0 + 8 * (j >> 1));
      }

      // ---- epilogue 2: accumulator 2 -> pre2 (HBM); the next chunk's tensor-memory load runs under the stores ----
      mbar_wait(bar_acc2f, lt & 1);
      tc_fence_after();
      {
        uint32_t r[32];
        tmem_ld32(lane_base + H + hf * 32, r);
#pragma unroll
        for (int kk = 0; kk < CPW; ++kk, ++nbox) {
          const int c0 = (hf + 2 * kk) * 32;
          const uint32_t buf = sbuf + (nbox & 1) * 2048;
          f32x2 bv[16];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 t4 = *reinterpret_cast<const float4*>(sb2 + c0 + 4 * q);
            bv[2 * q] = f2_make(t4.x, t4.y); bv[2 * q + 1] = f2_make(t4.z, t4.w);
          }
          if (OUT32) {
            // fp32 output: the chunk (+ bias) goes through the warp's staging as a swizzled 32x32 fp32 tile and is read
            // back row-major, so the residual load and the store are 512 contiguous bytes per instruction
            // residual rows of the whole chunk requested before anything waits (see tc_gemm.cu's fp32 epilogue)
            float4 pre[8];
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              const int i = it * 32 + lane, rr = i >> 3, ch = i & 7;
              pre[it] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (g.res && row0 + rr < g.M)
                pre[it] = *reinterpret_cast<const float4*>(g.res + (size_t)(row0 + rr) * g.ldr + c0 + ch * 4);
            }
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // epilogue 1's pre1 boxes
            __syncwarp();
            tmem_ld_wait();
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const f32x2 a0 = f2_add(f2_make_u(r[4 * q], r[4 * q + 1]), bv[2 * q]);
              const f32x2 a1 = f2_add(f2_make_u(r[4 * q + 2], r[4 * q + 3]), bv[2 * q + 1]);
              const uint32_t addr = sbuf + lane * 128 + ((q ^ (lane & 7)) << 4);
              asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "f"(f2_lo(a0)), "f"(f2_hi(a0)),
                           "f"(f2_lo(a1)), "f"(f2_hi(a1)) : "memory");
            }
            if (kk + 1 < CPW) {
              tmem_ld32(lane_base + H + c0 + 64, r);
            } else {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive_cluster(acc2e0);
            }
            __syncwarp();
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              const int i = it * 32 + lane, rr = i >> 3, ch = i & 7;
              const int64_t grow = row0 + rr;
              const int col = c0 + ch * 4;
              float4 v;
              asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                           : "r"(sbuf + rr * 128 + ((ch ^ (rr & 7)) << 4)) : "memory");
              if (grow < g.M) {
                if (g.res) { v.x += pre[it].x; v.y += pre[it].y; v.z += pre[it].z; v.w += pre[it].w; }
                *reinterpret_cast<float4*>(g.out2_f32 + (size_t)grow * g.ldo2f + col) = v;
                if (g.copy16)
                  *reinterpret_cast<uint2*>(g.copy16 + (size_t)grow * g.ldc16 + col) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
              }
            }
            __syncwarp();
            continue;
          }
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          tmem_ld_wait();
          uint32_t o[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) o[i] = f2_pack_bf16(f2_add(f2_make_u(r[2 * i], r[2 * i + 1]), bv[i]));
          if (kk + 1 < CPW) {
            tmem_ld32(lane_base + H + c0 + 64, r);
          } else {
            // accumulator 2 fully read by this warp: hand it back to the MMA warp before the stores
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
          if (lane == 0 && row0 < g.M) {
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                         :: "l"((uint64_t)&tmO2), "r"(c0), "r"((int)row0), "r"(buf) : "memory");
          }
          if (lane == 0) asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  __syncwarp();
  tc_fence_before();
  cluster_sync_all();            // nobody leaves (or frees tensor memory) while the peer can still signal it
  if (warp == kWarpMma) {
    tc_fence_after();
    tmem_dealloc_2cta(tmem_base, TMEM_COLS);
  }
}

template <int K1, int H, int TF1, bool SAVE, bool GATHER, bool OUT32, bool BOX = false>
int launch_chain(const TcChainArgs& g, const CUtensorMap& ta, const CUtensorMap& t1, const CUtensorMap& t2, const CUtensorMap& tg,
                 cudaStream_t st) {
  constexpr int smem = smem_bytes(K1, H);
  static_assert(smem <= kSmemMax, "shared memory budget");
  auto kern = tc_chain_kernel<K1, H, TF1, SAVE, GATHER, OUT32, BOX>;
  static int max_clusters = 0;                     // co-resident CTA pairs (one per TPC): 74 on a full B200
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
  kern<<<2 * ncl, kThreads, smem, st>>>(ta, t1, t2, tg, g);
  AE_LAUNCH_CHECK();
  return 0;
}

template <int K1, int H, int TF1>
int launch_tf(const TcChainArgs& g, const CUtensorMap& ta, const CUtensorMap& t1, const CUtensorMap& t2, const CUtensorMap& tg,
              bool save, cudaStream_t st) {
  const bool gather = g.gat16 != nullptr;
  if (g.out2_f32) {
    // node update: fp32 output with residual, no gather, plain (already activated / linear) operand
    if constexpr (K1 == H && TF1 == TC_TF_NONE) {
      if (gather) return AETHER_E_UNSUPPORTED;
      return save ? launch_chain<K1, H, TF1, true, false, true>(g, ta, t1, t2, tg, st)
                  : launch_chain<K1, H, TF1, false, false, true>(g, ta, t1, t2, tg, st);
    } else {
      return AETHER_E_UNSUPPORTED;
    }
  }
  if constexpr (K1 == H && TF1 == TC_TF_SILU) {
    if (gather && g.box_gather)
      return save ? launch_chain<K1, H, TF1, true, true, false, true>(g, ta, t1, t2, tg, st)
                  : launch_chain<K1, H, TF1, false, true, false, true>(g, ta, t1, t2, tg, st);
  }
  if (save) return gather ? launch_chain<K1, H, TF1, true, true, false>(g, ta, t1, t2, tg, st) : launch_chain<K1, H, TF1, true, false, false>(g, ta, t1, t2, tg, st);
  return gather ? launch_chain<K1, H, TF1, false, true, false>(g, ta, t1, t2, tg, st) : launch_chain<K1, H, TF1, false, false, false>(g, ta, t1, t2, tg, st);
}

template <int K1, int H>
int launch_kh(const TcChainArgs& g, const CUtensorMap& ta, const CUtensorMap& t1, const CUtensorMap& t2, const CUtensorMap& tg,
              int tf1, bool save, cudaStream_t st) {
  if (tf1 == TC_TF_SILU) return launch_tf<K1, H, TC_TF_SILU>(g, ta, t1, t2, tg, save, st);
  return launch_tf<K1, H, TC_TF_NONE>(g, ta, t1, t2, tg, save, st);
}

}  // namespace

bool tc_chain_supported(int K1, int H) {
  return (H == 64 || H == 128 || H == 256) && (K1 == 64 || K1 == H);
}

int tc_chain(const TcChainArgs& g, const void* A, int64_t lda, int K1, int H, int tf1, cudaStream_t st) {
  if (g.M <= 0) return 0;
  if (!tc_chain_supported(K1, H)) return AETHER_E_UNSUPPORTED;
  if (!g.wpack1 || !g.wpack2 || ((g.out2 != nullptr) == (g.out2_f32 != nullptr))) return AETHER_E_BADARG;
  if (g.out2_f32 && ((g.ldo2f & 3) || (g.res && (g.ldr & 3)) || (g.copy16 && (g.ldc16 & 3)))) return AETHER_E_BADARG;
  if (g.gat16 && (!g.idx_a || !g.idx_b)) return AETHER_E_BADARG;
  if (g.gat16 && (((uintptr_t)g.gat16 & 15) || (g.ldg & 7) || (g.goff_b & 7))) return AETHER_E_BADARG;   // 128-bit row loads
  const bool save = g.out1 != nullptr;
  CUtensorMap ta, t1, t2, tg;
  memset(&t1, 0, sizeof(t1));
  memset(&tg, 0, sizeof(tg));
  if (g.gat16 && g.box_gather) {
    if (g.gat_rows <= 0) return AETHER_E_BADARG;
    AE_CHECK(tc_make_tmap_bf16(&tg, g.gat16, g.gat_rows, H, g.ldg, kTileM));
  }
  AE_CHECK(tc_make_tmap_bf16(&ta, A, g.M, K1, lda, kTileM));
  if (save) AE_CHECK(tc_make_tmap_bf16(&t1, g.out1, g.M, H, g.ldo1, 32, 32));
  memset(&t2, 0, sizeof(t2));
  if (g.out2) AE_CHECK(tc_make_tmap_bf16(&t2, g.out2, g.M, H, g.ldo2, 32, 32));
#define AE_CH_CASE(KK, HH) if (K1 == KK && H == HH) return launch_kh<KK, HH>(g, ta, t1, t2, tg, tf1, save, st);
  AE_CH_CASE(256, 256) AE_CH_CASE(64, 256) AE_CH_CASE(128, 128) AE_CH_CASE(64, 128) AE_CH_CASE(64, 64)
#undef AE_CH_CASE
  return AETHER_E_UNSUPPORTED;
}

}  // namespace aether
