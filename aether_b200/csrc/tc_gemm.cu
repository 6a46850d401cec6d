// Persistent tcgen05 GEMM over edge (or node) rows:
//
//     out[M, N] = epi( tf(A[M, K]) * B^T )          A bf16 row-major in HBM, B = packed bf16 weights
//
// This is every Linear of GNNLayer.message_fn (nn/state2state/locs/locs.py:206-212,235) in the bf16
// path and the data-gradient GEMMs of its backward.  One CTA per SM, warp-specialised:
//
//   warp 8      TMA producer : the weight image once (cp.async.bulk), then one 128x64 bf16 box of A
//               per pipeline slot (cp.async.bulk.tensor, 128-byte swizzle = the UMMA K-major layout)
//   warps 10-13 transform    : optional in-place SiLU on the landed box (the consumer of a stored
//               pre-activation applies the activation, so activations are never written to HBM)
//   warp 9      MMA issuer   : one thread, tcgen05.mma 128 x N x 16 per instruction, fp32
//               accumulators in TMEM, two accumulators ping-pong between MMA and epilogue
//   warps 0-7   epilogue     : two warps per TMEM lane quarter (warp & 3), alternating 32-column chunks;
//               tcgen05.ld (one output row per thread), + bias, + gathered node terms
//               P_s[send] + P_r[recv] (node-factorised first Linear), x silu'(saved pre-activation)
//               for the backward chain, bf16 (32x32 boxes staged in shared memory, TMA store) or fp32 store.
//               Two warps per scheduler hide each other's TMEM / global-load latency.
//
// Shared memory: weights N*K*2 bytes resident for the whole kernel (128 KB at N=K=256), the rest is
// the A ring (16 KB per slot, 4 slots at N=K=256) and 32 KB of epilogue staging.  HBM traffic per row: 2K bytes in, 2N bytes out.
#include <string.h>

#include <type_traits>

#include "tc_gemm.cuh"

namespace aether {

namespace {

using namespace tc;

#ifdef AETHER_TC_REGSPLIT
// EXPERIMENTAL, off by default and not measured yet: 16 warps so that every warpgroup is complete; after
// the set-up barrier the two epilogue warpgroups raise their register budget with setmaxnreg and the producer / MMA /
// transform warpgroups give theirs up (8 x 32 x (128 - 56) registers change hands; warps 14-15 idle).  The larger budget
// holds FOUR operand slots, so gathered rows (and the silu' rows that go with them) are requested two chunks ahead.
constexpr int kThreads = 512;
constexpr int kRegsEpi = 200, kRegsAux = 56;
#else
constexpr int kThreads = 448;
#endif
constexpr int kEpiWarps = 8;
constexpr int kWarpTma = 8, kWarpMma = 9, kWarpTf0 = 10;
constexpr int kTileM = 128;
constexpr int kSlotBytes = kTileM * 128;          // 128 rows x 64 bf16
constexpr int kSmemMax = 232448;                  // 227 KB
constexpr int kMiscBytes = 2048;                  // barriers + tmem pointer + bias

__host__ __device__ constexpr int weight_bytes(int K, int N) { return K * N * 2; }
constexpr int kStageBytes = kEpiWarps * 2 * 2048; // per epilogue warp: two 32x32 bf16 boxes, or one 32x32 fp32 tile
__host__ __device__ constexpr int num_slots(int K, int N, bool staged) {
  (void)staged;
  int s = (kSmemMax - 1024 - kMiscBytes - weight_bytes(K, N) - kStageBytes) / kSlotBytes;
#ifdef AETHER_TC_MAX_SLOTS                          // experiment knob: ring-depth sensitivity (e.g. -DAETHER_TC_MAX_SLOTS=3)
  if (s > AETHER_TC_MAX_SLOTS) s = AETHER_TC_MAX_SLOTS;
#endif
  return s > 8 ? 8 : s;
}
__host__ __device__ constexpr int smem_bytes(int K, int N, bool staged) {
  return 1024 + weight_bytes(K, N) + num_slots(K, N, staged) * kSlotBytes + kStageBytes + kMiscBytes;
}
__host__ __device__ constexpr int tmem_cols(int N) { return 2 * N < 32 ? 32 : 2 * N; }

template <int K, int N, int TF, bool STAGED>
__global__ void __launch_bounds__(kThreads, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmO, TcGemmArgs g) {
  static_assert(K % 64 == 0 && N % 32 == 0 && N <= 256 && (N & (N - 1)) == 0, "unsupported shape");
  static_assert(!STAGED || N % 64 == 0, "staged epilogue works on 64-column boxes");
  constexpr int KC = K / 64;                      // 64-wide K chunks = pipeline slots per tile
  constexpr int S = num_slots(K, N, STAGED);
  constexpr uint32_t IDESC = make_idesc(128, N, MAJOR_K, MAJOR_K);
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sW = base;
  const uint32_t sA = base + weight_bytes(K, N);
  const uint32_t sStage = sA + S * kSlotBytes;
  const uint32_t sMisc = sStage + kStageBytes;
  // barriers (8 bytes each)
  const uint32_t bar_raw = sMisc;                 // [S] TMA landed
  const uint32_t bar_rdy = sMisc + 64;            // [S] transform done
  const uint32_t bar_free = sMisc + 128;          // [S] MMA consumed
  const uint32_t bar_accf = sMisc + 192;          // [2] accumulator full
  const uint32_t bar_acce = sMisc + 208;          // [2] accumulator drained
  const uint32_t bar_w = sMisc + 224;             // weights landed
  const uint32_t tmem_slot = sMisc + 240;
  uint8_t* misc_ptr = smem_raw + (sMisc - smem_u32(smem_raw));
  float* sbias = reinterpret_cast<float*>(misc_ptr + 1024);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t n_tiles = (g.M + kTileM - 1) / kTileM;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(bar_raw + 8 * s, 1);
      mbar_init(bar_rdy + 8 * s, 128);
      mbar_init(bar_free + 8 * s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(bar_accf + 8 * a, 1);
      mbar_init(bar_acce + 8 * a, kEpiWarps);
    }
    mbar_init(bar_w, 1);
    fence_mbar_init();
  }
  if (threadIdx.x < N) sbias[threadIdx.x] = g.bias ? g.bias[threadIdx.x] : 0.0f;
  if (warp == kWarpMma) tmem_alloc(tmem_slot, tmem_cols(N));
  if (warp == kWarpTma && lane == 0) { tma_prefetch_desc(&tmA); if (STAGED) tma_prefetch_desc(&tmO); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
#ifdef AETHER_TC_REGSPLIT
  // one instruction for both auxiliary warpgroups; the epilogue warpgroups raise theirs inside their own branch, so
  // ptxas allocates that branch against the larger budget
  if (warp >= kEpiWarps) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" :: "n"(kRegsAux));
#endif

  if (warp == kWarpTma) {
    // ===== TMA producer =====
    if (lane == 0) {
      mbar_arrive_expect_tx(bar_w, weight_bytes(K, N));
      for (int c = 0; c < KC; ++c)
        bulk_load(sW + c * (N * 128), (const uint8_t*)g.wpack + (size_t)c * (N * 128), N * 128, bar_w);
      uint32_t it = 0;
      for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        for (int c = 0; c < KC; ++c, ++it) {
          const uint32_t s = it % S, ph = (it / S) & 1;
          mbar_wait(bar_free + 8 * s, ph ^ 1);
          mbar_arrive_expect_tx(bar_raw + 8 * s, kSlotBytes);
          tma_load_2d(sA + s * kSlotBytes, &tmA, c * 64, (int)(t * kTileM), bar_raw + 8 * s);
        }
      }
    }
  } else if (warp == kWarpMma) {
    // ===== MMA issuer =====
    if (lane == 0) {
      mbar_wait(bar_w, 0);
      uint32_t it = 0, lt = 0;
      for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++lt) {
        const uint32_t acc = lt & 1;
        mbar_wait(bar_acce + 8 * acc, ((lt >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d = tmem_base + acc * N;
        for (int c = 0; c < KC; ++c, ++it) {
          const uint32_t s = it % S, ph = (it / S) & 1;
          mbar_wait((TF != TC_TF_NONE ? bar_rdy : bar_raw) + 8 * s, ph);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t ad = make_sdesc(sA + s * kSlotBytes + k * 32, 16, 1024);
            const uint64_t bd = make_sdesc(sW + c * (N * 128) + k * 32, 16, 1024);
            umma_bf16(d, ad, bd, IDESC, (c | k) != 0);
          }
          umma_commit(bar_free + 8 * s);
        }
        umma_commit(bar_accf + 8 * acc);
      }
    }
  } else if (warp >= kWarpTf0) {
    // ===== transform: in-place SiLU on the landed A box =====
#ifdef AETHER_TC_REGSPLIT
    if (warp >= kWarpTf0 + 4) {
      // warps 14-15 only complete the last warpgroup
    } else
#endif
    if (TF != TC_TF_NONE) {
      const int tt = threadIdx.x - kWarpTf0 * 32; // 0..127
      uint32_t it = 0;
      for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        for (int c = 0; c < KC; ++c, ++it) {
          const uint32_t s = it % S, ph = (it / S) & 1;
          mbar_wait(bar_raw + 8 * s, ph);
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
          mbar_arrive(bar_rdy + 8 * s);
        }
      }
    }
  } else if (STAGED) {
    // ===== epilogue, bf16 output: TMEM -> registers (one row per thread) -> swizzled shared staging
    //       -> TMA store; gathered node terms and the silu' factor are applied in registers =====
    // Each warp owns two 32-row x 32-column staging boxes (2 KB each, 64-byte swizzle), so only
    // __syncwarp is needed.  Warps w and w+4 share a lane quarter and take alternate column chunks.
    // The loop below exists twice: ZDEEP = the epilogue's only operand is the silu' rows (no gathers), which then are
    // requested TWO chunks ahead (three register slots); otherwise one chunk ahead (gathers + silu' share the register
    // budget: the two-ahead variant with a gathered row spilled and measured 30 % slower).
    // GDEEP (AETHER_TC_REGSPLIT builds only): gathered rows and their companions two chunks ahead, four slots.
#ifdef AETHER_TC_REGSPLIT
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" :: "n"(kRegsEpi));
#endif
    auto run = [&](auto zdeep_tag, auto gdeep_tag) {
      constexpr bool ZDEEP = decltype(zdeep_tag)::value;
      constexpr bool GDEEP = decltype(gdeep_tag)::value;
#ifdef AETHER_TC_REGSPLIT
      constexpr bool ZD3 = ZDEEP;                   // with the larger budget the silu' rows go THREE chunks ahead (slot x2)
#else
      constexpr bool ZD3 = false;
#endif
      const int qr = warp & 3, hf = warp >> 2;
      const uint32_t sbuf = sStage + warp * 4096;
      uint32_t lt = 0, nbox = 0;
      const bool has_b = g.gat16 && g.idx_b;      // tc_gemm() rejects idx_b together with dsilu: they share a slot
      // Every epilogue operand of a chunk (64 contiguous bytes per row and source) is requested ONE CHUNK AHEAD,
      // also across tiles, so its L2 / HBM latency is covered by the previous chunk's work.  Two register slots:
      // xn = gathered row idx_a, yn = gathered row idx_b or the silu' rows.
      uint4 xn[4], yn[4], y2[4], x2[4];
      int ia_n = 0, ib_n = 0;
      constexpr int CPT = N >= 64 ? N / 64 : 1;     // chunks per tile and warp
      // ZDEEP: silu' rows of the chunk d chunks after (tile row `row_`, chunk k_ of its tile) into dst
      auto zrequest = [&](uint4 (&dst)[4], int64_t t_, int64_t row_, int k_) {
        const int m = k_ / CPT, kk = k_ % CPT;
        if (t_ + (int64_t)m * gridDim.x >= n_tiles) return;
        int64_t rr = row_ + (int64_t)m * gridDim.x * kTileM;
        rr = rr < g.M ? rr : g.M - 1;
        const uint4* pd = reinterpret_cast<const uint4*>(g.dsilu + (size_t)rr * g.ldd + hf * 32 + kk * 64);
#pragma unroll
        for (int q = 0; q < 4; ++q) dst[q] = __ldg(pd + q);
      };
      auto request = [&](uint4 (&dx)[4], uint4 (&dy)[4], int ia_, int ib_, int64_t zrow_, int c) {
        if (g.gat16) {
          const uint4* pa = reinterpret_cast<const uint4*>(g.gat16 + (size_t)ia_ * g.ldg + c);
#pragma unroll
          for (int q = 0; q < 4; ++q) dx[q] = __ldg(pa + q);
        }
        if (has_b) {
          const uint4* pb = reinterpret_cast<const uint4*>(g.gat16 + ((int64_t)ib_ * g.ldg + g.goff_b + c));
#pragma unroll
          for (int q = 0; q < 4; ++q) dy[q] = __ldg(pb + q);
        } else if (g.dsilu) {
          const uint4* pd = reinterpret_cast<const uint4*>(g.dsilu + (size_t)zrow_ * g.ldd + c);
#pragma unroll
          for (int q = 0; q < 4; ++q) dy[q] = __ldg(pd + q);
        }
      };
      if ((int64_t)blockIdx.x < n_tiles) {
        const int64_t r0 = (int64_t)blockIdx.x * kTileM + qr * 32 + lane;
        if constexpr (ZDEEP) {
          zrequest(yn, blockIdx.x, r0, 0);
          zrequest(y2, blockIdx.x, r0, 1);
          if constexpr (ZD3) zrequest(x2, blockIdx.x, r0, 2);
        } else {
          if (g.gat16 && r0 < g.M) { ia_n = g.idx_a[r0]; ib_n = g.idx_b ? g.idx_b[r0] : 0; }
          request(xn, yn, ia_n, ib_n, r0 < g.M ? r0 : g.M - 1, hf * 32);
          if constexpr (GDEEP) request(x2, y2, ia_n, ib_n, r0 < g.M ? r0 : g.M - 1, hf * 32 + 64);
        }
      }
      for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++lt) {
        const uint32_t acc = lt & 1;
        const int64_t row0 = t * kTileM + qr * 32;
        const int64_t row = row0 + lane;
        const int ia = ia_n, ib = ib_n;
        const int64_t zrow = row < g.M ? row : g.M - 1;     // rows past the end read valid memory, TMA clips them
        const int64_t tn = t + gridDim.x;
        const int64_t row_n = row + (int64_t)gridDim.x * kTileM;
        const int64_t zrow_n = row_n < g.M ? row_n : g.M - 1;
        ia_n = 0; ib_n = 0;
        if (g.gat16 && tn < n_tiles && row_n < g.M) { ia_n = g.idx_a[row_n]; ib_n = g.idx_b ? g.idx_b[row_n] : 0; }
        mbar_wait(bar_accf + 8 * acc, (lt >> 1) & 1);
        tc_fence_after();
#pragma unroll 1
        for (int j = hf; j < N / 32; j += 2, ++nbox) {
          const uint32_t buf = sbuf + (nbox & 1) * 2048;
          const int c0 = j * 32;
          uint4 a4[4], z4[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) { a4[q] = xn[q]; z4[q] = yn[q]; }
          if constexpr (ZDEEP) {
#pragma unroll
            for (int q = 0; q < 4; ++q) yn[q] = y2[q];
            if constexpr (ZD3) {
#pragma unroll
              for (int q = 0; q < 4; ++q) y2[q] = x2[q];
              zrequest(x2, t, row, (j >> 1) + 3);
            } else {
              zrequest(y2, t, row, (j >> 1) + 2);
            }
          } else if constexpr (GDEEP) {
#pragma unroll
            for (int q = 0; q < 4; ++q) { xn[q] = x2[q]; yn[q] = y2[q]; }
            const int k2 = (j >> 1) + 2;            // this warp's chunk two ahead: same tile, or the next one
            if (k2 < CPT) request(x2, y2, ia, ib, zrow, c0 + 128);
            else if (tn < n_tiles) request(x2, y2, ia_n, ib_n, zrow_n, hf * 32 + (k2 - CPT) * 64);
          } else {
            if (j + 2 < N / 32) request(xn, yn, ia, ib, zrow, c0 + 64);
            else if (tn < n_tiles) request(xn, yn, ia_n, ib_n, zrow_n, hf * 32);
          }
          uint32_t r[32];
          tmem_ld32(tmem_base + ((uint32_t)(qr * 32) << 16) + acc * N + c0, r);
          // the TMA store that read this buffer two boxes ago must have finished reading it
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          tmem_ld_wait();
          if (j + 2 >= N / 32) {
            // accumulator fully read by this warp: hand it back to the MMA warp before the slow part
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_acce + 8 * acc);
          }
          float v[32];
          if (g.bias) {
            const float* bs = sbias + c0;
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) + bs[i];
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
          }
          if (has_b) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint32_t bb[4] = {z4[q].x, z4[q].y, z4[q].z, z4[q].w};
#pragma unroll
              for (int e = 0; e < 4; ++e) { v[8 * q + 2 * e] += bf16_lo(bb[e]); v[8 * q + 2 * e + 1] += bf16_hi(bb[e]); }
            }
          }
          if (g.dsilu) {
            // backward chain: out = (acc [+ gathered term]) * silu'(z).  w = 2 silu'(z) - 1 = t + h (1 - t^2)
            // (h = z/2, t = tanh h, |w| <= 1.2) is evaluated two elements per instruction in bf16x2; the product
            // with the fp32 accumulator stays in fp32: out = v/2 + (v/2) w.
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint32_t zz[4] = {z4[q].x, z4[q].y, z4[q].z, z4[q].w};
              const uint32_t aa[4] = {a4[q].x, a4[q].y, a4[q].z, a4[q].w};
              uint32_t o[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const uint32_t w = bf2_dsilu_w(zz[e]);
                float v0 = v[8 * q + 2 * e], v1 = v[8 * q + 2 * e + 1];
                if (g.gat16) { v0 += bf16_lo(aa[e]); v1 += bf16_hi(aa[e]); }
                v0 *= 0.5f; v1 *= 0.5f;
                o[e] = pack_bf16(fmaf(v0, bf16_lo(w), v0), fmaf(v1, bf16_hi(w), v1));
              }
              const uint32_t addr = buf + lane * 64 + ((q ^ ((lane >> 1) & 3)) << 4);
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
            }
          } else {
            if (g.gat16) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const uint32_t aa[4] = {a4[q].x, a4[q].y, a4[q].z, a4[q].w};
#pragma unroll
                for (int e = 0; e < 4; ++e) { v[8 * q + 2 * e] += bf16_lo(aa[e]); v[8 * q + 2 * e + 1] += bf16_hi(aa[e]); }
              }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              uint32_t o[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) o[e] = pack_bf16(v[8 * q + 2 * e], v[8 * q + 2 * e + 1]);
              const uint32_t addr = buf + lane * 64 + ((q ^ ((lane >> 1) & 3)) << 4);
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0 && row0 < g.M) {
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                         :: "l"((uint64_t)&tmO), "r"(c0), "r"((int)row0), "r"(buf) : "memory");
          }
          if (lane == 0) asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
    };
    bool done = false;
#ifdef AETHER_TC_REGSPLIT
    if constexpr (N >= 256) {
      if (g.gat16) { run(std::false_type{}, std::true_type{}); done = true; }
    }
#endif
    if (done) {
    } else if (g.dsilu && !g.gat16) run(std::true_type{}, std::false_type{});
    else run(std::false_type{}, std::false_type{});
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  } else {
    // ===== epilogue, fp32 output (node-level Linears, layer-1 feature gradients) =====
    // Row-per-thread global accesses would touch 32 lines per instruction; instead the accumulator chunk (+ bias) goes
    // through the warp's swizzled 32x32 fp32 staging tile and is read back 16 bytes per lane in row-major order, so
    // the residual / silu' / mask / accumulate operands and the store are all 512 contiguous bytes per instruction.
#ifdef AETHER_TC_REGSPLIT
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" :: "n"(kRegsEpi));
#endif
    const uint32_t sbuf = sStage + warp * 4096;
    const bool vec = (g.ldo32 & 3) == 0 && (g.n_valid & 3) == 0 && (!g.res || (g.ldr & 3) == 0) &&
                     (!g.dsilu32 || (g.ldd32 & 3) == 0) && (!g.emask || (g.ldm & 3) == 0);
    uint32_t lt = 0;
    for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++lt) {
      const uint32_t acc = lt & 1;
      mbar_wait(bar_accf + 8 * acc, (lt >> 1) & 1);
      tc_fence_after();
      const int qr = warp & 3;
      const int64_t row = t * kTileM + qr * 32 + lane;
      const bool live = row < g.M;
      if (vec) {
#pragma unroll 1
        for (int j = warp >> 2; j < N / 32; j += 2) {
          uint32_t r[32];
          tmem_ld32(tmem_base + ((uint32_t)(qr * 32) << 16) + acc * N + j * 32, r);
          // the residual / accumulate operands of the whole chunk are requested before anything waits (they used to be
          // loaded one by one right in front of their use: 8 exposed global-load latencies per chunk, 50 k cycles per
          // tile -- profiles/r02_ncu_node_gemm.md)
          float4 pre[8];
          const float* psrc = g.res ? g.res : (g.accumulate ? g.out_f32 : nullptr);
          const int64_t pld = g.res ? g.ldr : g.ldo32;
          const bool both = g.res && g.accumulate;
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int i = it * 32 + lane, rr = i >> 3, ch = i & 7;
            const int64_t grow = t * kTileM + qr * 32 + rr;
            const int col = j * 32 + ch * 4;
            pre[it] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (psrc && grow < g.M && col < g.n_valid) pre[it] = *reinterpret_cast<const float4*>(psrc + (size_t)grow * pld + col);
          }
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 b4 = *reinterpret_cast<const float4*>(sbias + j * 32 + 4 * q);
            const uint32_t addr = sbuf + lane * 128 + ((q ^ (lane & 7)) << 4);
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "f"(__uint_as_float(r[4 * q]) + b4.x),
                         "f"(__uint_as_float(r[4 * q + 1]) + b4.y), "f"(__uint_as_float(r[4 * q + 2]) + b4.z),
                         "f"(__uint_as_float(r[4 * q + 3]) + b4.w) : "memory");
          }
          __syncwarp();
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int i = it * 32 + lane, rr = i >> 3, ch = i & 7;
            const int64_t grow = t * kTileM + qr * 32 + rr;
            const int col = j * 32 + ch * 4;
            float4 v;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                         : "r"(sbuf + rr * 128 + ((ch ^ (rr & 7)) << 4)) : "memory");
            if (grow < g.M && col < g.n_valid) {
              if (g.res) { v.x += pre[it].x; v.y += pre[it].y; v.z += pre[it].z; v.w += pre[it].w; }
              if (g.dsilu32) {
                const float4 z = __ldg(reinterpret_cast<const float4*>(g.dsilu32 + (size_t)grow * g.ldd32 + col));
                v.x *= dsilu_f(z.x); v.y *= dsilu_f(z.y); v.z *= dsilu_f(z.z); v.w *= dsilu_f(z.w);
              }
              if (g.emask) {
                const float4 z = __ldg(reinterpret_cast<const float4*>(g.emask + (size_t)grow * g.ldm + col));
                v.x *= z.x; v.y *= z.y; v.z *= z.z; v.w *= z.w;
              }
              float4* po = reinterpret_cast<float4*>(g.out_f32 + (size_t)grow * g.ldo32 + col);
              if (g.accumulate) {
                const float4 a = both ? *po : pre[it];
                v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
              }
              *po = v;
            }
          }
          __syncwarp();
        }
      } else {
#pragma unroll 1
      for (int j = warp >> 2; j < N / 32; j += 2) {
        uint32_t r[32];
        tmem_ld32(tmem_base + ((uint32_t)(qr * 32) << 16) + acc * N + j * 32, r);
        tmem_ld_wait();
        if (!live) continue;
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) + sbias[j * 32 + i];
        if (g.res) {
          const float4* pr = reinterpret_cast<const float4*>(g.res + (size_t)row * g.ldr + j * 32);
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 a = pr[q];
            v[4 * q] += a.x; v[4 * q + 1] += a.y; v[4 * q + 2] += a.z; v[4 * q + 3] += a.w;
          }
        }
        if (g.dsilu32) {
          const float4* pz = reinterpret_cast<const float4*>(g.dsilu32 + (size_t)row * g.ldd32 + j * 32);
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 z = __ldg(pz + q);
            v[4 * q] *= dsilu_f(z.x); v[4 * q + 1] *= dsilu_f(z.y); v[4 * q + 2] *= dsilu_f(z.z); v[4 * q + 3] *= dsilu_f(z.w);
          }
        }
        if (g.emask) {
          const float4* pm = reinterpret_cast<const float4*>(g.emask + (size_t)row * g.ldm + j * 32);
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 z = __ldg(pm + q);
            v[4 * q] *= z.x; v[4 * q + 1] *= z.y; v[4 * q + 2] *= z.z; v[4 * q + 3] *= z.w;
          }
        }
        if (g.accumulate) {
          const float4* pc = reinterpret_cast<const float4*>(g.out_f32 + (size_t)row * g.ldo32 + j * 32);
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 a = pc[q];
            v[4 * q] += a.x; v[4 * q + 1] += a.y; v[4 * q + 2] += a.z; v[4 * q + 3] += a.w;
          }
        }
        float* po = g.out_f32 + (size_t)row * g.ldo32 + j * 32;
        if (j * 32 + 32 <= g.n_valid && (g.ldo32 & 3) == 0) {
#pragma unroll
          for (int q = 0; q < 8; ++q)
            reinterpret_cast<float4*>(po)[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (j * 32 + i < g.n_valid) po[i] = v[i];
        }
      }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_acce + 8 * acc);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kWarpMma) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols(N));
  }
}

// fp32 weights -> bf16 shared-memory image of the B operand (K-major, 128-byte swizzle):
// logical B[n][k] = transpose ? W[k*ldw + n] : W[n*ldw + (kmod ? k % kmod : k)], zero outside
// [n_valid, k_valid].
__global__ void pack_weight_kernel(const float* __restrict__ W, int64_t ldw, int N, int K, int n_valid, int k_valid,
                                   int transpose, int kmod, uint8_t* __restrict__ out) {
  const int u = blockIdx.x * blockDim.x + threadIdx.x;   // one 16-byte unit = 8 consecutive k
  if (u >= N * K / 8) return;
  const int n = u % N, kb = u / N;                       // kb: 8-element block index along K
  uint32_t w[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    float f[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int k = kb * 8 + 2 * e + h;
      const int ks = kmod ? k % kmod : k;
      float x = 0.0f;
      if (n < n_valid && ks < k_valid) x = transpose ? W[(size_t)ks * ldw + n] : W[(size_t)n * ldw + ks];
      f[h] = x;
    }
    w[e] = tc::pack_bf16(f[0], f[1]);
  }
  const int c = kb / 8, j = kb % 8;
  const size_t off = (size_t)c * N * 128 + (size_t)n * 128 + (size_t)((j ^ (n & 7)) * 16);
  *reinterpret_cast<uint4*>(out + off) = make_uint4(w[0], w[1], w[2], w[3]);
}

struct PackTable {
  TcPackJob job[TC_PACK_MAX_JOBS];
  int32_t unit_end[TC_PACK_MAX_JOBS];     // exclusive prefix end of each job's 16-byte units
  int32_t n;
};

__global__ void pack_weights_batched_kernel(const __grid_constant__ PackTable t) {
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= t.unit_end[t.n - 1]) return;
  int lo = 0, hi = t.n - 1;                // first job whose end is past u
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (u < t.unit_end[mid]) hi = mid; else lo = mid + 1;
  }
  const TcPackJob& j = t.job[lo];
  const int lu = u - (lo ? t.unit_end[lo - 1] : 0);
  const int n = lu % j.N, kb = lu / j.N;
  uint32_t w[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    float f[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int k = kb * 8 + 2 * e + h;
      const int ks = j.kmod ? k % j.kmod : k;
      float x = 0.0f;
      if (n < j.n_valid && ks < j.k_valid) x = j.transpose ? j.W[(size_t)ks * j.ldw + n] : j.W[(size_t)n * j.ldw + ks];
      f[h] = x;
    }
    w[e] = tc::pack_bf16(f[0], f[1]);
  }
  const int c = kb / 8, jj = kb % 8;
  const size_t off = (size_t)c * j.N * 128 + (size_t)n * 128 + (size_t)((jj ^ (n & 7)) * 16);
  *reinterpret_cast<uint4*>((uint8_t*)j.out + off) = make_uint4(w[0], w[1], w[2], w[3]);
}

__global__ void convert_bf16_kernel(const float* __restrict__ A, int64_t lda, int64_t M, int cols, int tf,
                                    const float* __restrict__ mask, __nv_bfloat16* __restrict__ out, int64_t ldo) {
  const int cq = cols / 4;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M * cq) return;
  const int64_t m = i / cq;
  const int c = (int)(i % cq) * 4;
  float4 v = *reinterpret_cast<const float4*>(A + (size_t)m * lda + c);
  if (tf != TF_NONE) {
    v.x = silu_f(v.x); v.y = silu_f(v.y); v.z = silu_f(v.z); v.w = silu_f(v.w);
    if (tf == TF_SILU_MASK) {
      const float4 k = *reinterpret_cast<const float4*>(mask + (size_t)m * lda + c);
      v.x *= k.x; v.y *= k.y; v.z *= k.z; v.w *= k.w;
    }
  }
  uint2 o;
  o.x = pack_bf16(v.x, v.y);
  o.y = pack_bf16(v.z, v.w);
  *reinterpret_cast<uint2*>(out + (size_t)m * ldo + c) = o;
}

template <int K, int N, int TF, bool STAGED>
int launch_one(const TcGemmArgs& g, const CUtensorMap& tm, const CUtensorMap& to, cudaStream_t st) {
  const int64_t tiles = (g.M + kTileM - 1) / kTileM;
  const int sms = tc_sm_count();
  const unsigned grid = (unsigned)(tiles < sms ? tiles : sms);
  constexpr int smem = smem_bytes(K, N, STAGED);
  static_assert(num_slots(K, N, STAGED) >= 2, "pipeline too shallow");
  cudaError_t e = cudaFuncSetAttribute(tc_gemm_kernel<K, N, TF, STAGED>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return (int)e;
  tc_gemm_kernel<K, N, TF, STAGED><<<grid, kThreads, smem, st>>>(tm, to, g);
  AE_LAUNCH_CHECK();
  return 0;
}

template <int K, int N>
int launch_kn(const TcGemmArgs& g, const CUtensorMap& tm, const CUtensorMap& to, int tf, bool staged, cudaStream_t st) {
  if constexpr (N >= 64) {
    if (staged) {
      if (tf == TC_TF_SILU) return launch_one<K, N, TC_TF_SILU, true>(g, tm, to, st);
      return launch_one<K, N, TC_TF_NONE, true>(g, tm, to, st);
    }
  }
  if (tf == TC_TF_SILU) return launch_one<K, N, TC_TF_SILU, false>(g, tm, to, st);
  return launch_one<K, N, TC_TF_NONE, false>(g, tm, to, st);
}

}  // namespace

int tc_sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    sms = n;
  }
  return sms;
}

int tc_make_tmap_bf16(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows,
                      int box_cols) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess) return (int)e;
    if (!p || q != cudaDriverEntryPointSuccess) return AETHER_E_UNSUPPORTED;
    fn = (EncodeFn)p;
  }
  if (rows <= 0 || (box_cols != 64 && box_cols != 32) || cols < box_cols || (ld * 2) % 16 || ((uintptr_t)base & 15))
    return AETHER_E_BADARG;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, box_cols == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : AETHER_E_BADARG;
}

size_t tc_weight_pack_bytes(int K, int N) { return (size_t)K * N * 2; }

int tc_pack_weight(const float* W, int64_t ldw, int N, int K, int n_valid, int k_valid, int transpose, int kmod,
                   void* out, cudaStream_t st) {
  if (K % 64 || N % 8 || ((uintptr_t)out & 15)) return AETHER_E_BADARG;
  const int units = N * K / 8;
  pack_weight_kernel<<<(units + 255) / 256, 256, 0, st>>>(W, ldw, N, K, n_valid, k_valid, transpose, kmod, (uint8_t*)out);
  AE_LAUNCH_CHECK();
  return 0;
}

int tc_pack_weights_batched(const TcPackJob* jobs, int n_jobs, cudaStream_t st) {
  if (n_jobs <= 0) return 0;
  if (n_jobs > TC_PACK_MAX_JOBS) return AETHER_E_BADARG;
  PackTable t;
  memset(&t, 0, sizeof(t));
  int tot = 0;
  for (int i = 0; i < n_jobs; ++i) {
    const TcPackJob& j = jobs[i];
    if (j.K % 64 || j.N % 8 || ((uintptr_t)j.out & 15) || !j.W) return AETHER_E_BADARG;
    t.job[i] = j;
    tot += j.N * j.K / 8;
    t.unit_end[i] = tot;
  }
  t.n = n_jobs;
  pack_weights_batched_kernel<<<(tot + 255) / 256, 256, 0, st>>>(t);
  AE_LAUNCH_CHECK();
  return 0;
}

int tc_convert_bf16(const float* A, int64_t lda, int64_t M, int cols, int tf, const float* mask, void* out, int64_t ldo,
                    cudaStream_t st) {
  if (M <= 0 || cols <= 0) return 0;
  if (cols % 4 || lda % 4 || ldo % 4 || ((uintptr_t)A & 15)) return AETHER_E_BADARG;
  const int64_t tot = M * (cols / 4);
  convert_bf16_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(A, lda, M, cols, tf, mask, (__nv_bfloat16*)out, ldo);
  AE_LAUNCH_CHECK();
  return 0;
}

bool tc_gemm_supported(int K, int N) {
  return (K == 64 || K == 128 || K == 256) && (N == 32 || N == 64 || N == 128 || N == 256);
}

int tc_gemm(const TcGemmArgs& g, const void* A, int64_t lda, int K, int N, int tf, cudaStream_t st) {
  if (g.M <= 0) return 0;
  if (!tc_gemm_supported(K, N)) return AETHER_E_UNSUPPORTED;
  // exactly one output: bf16 (staged TMA-store epilogue) or fp32 (direct stores)
  const bool staged = g.out_bf16 != nullptr;
  if (staged == (g.out_f32 != nullptr)) return AETHER_E_BADARG;
  if (staged && (N < 64 || g.res || g.dsilu32 || g.emask || g.accumulate)) return AETHER_E_BADARG;
  if (!staged && (g.gat16 || g.dsilu)) return AETHER_E_BADARG;
  if (g.idx_b && g.dsilu) return AETHER_E_BADARG;   // the second gathered row and the silu' rows share a register slot
  CUtensorMap tm, to;
  memset(&to, 0, sizeof(to));
  AE_CHECK(tc_make_tmap_bf16(&tm, A, g.M, K, lda, kTileM));
  if (staged) AE_CHECK(tc_make_tmap_bf16(&to, g.out_bf16, g.M, N, g.ldo, 32, 32));
#define AE_TC_CASE(KK, NN) if (K == KK && N == NN) return launch_kn<KK, NN>(g, tm, to, tf, staged, st);
  AE_TC_CASE(256, 256) AE_TC_CASE(256, 32) AE_TC_CASE(64, 256)
  AE_TC_CASE(128, 128) AE_TC_CASE(128, 32) AE_TC_CASE(64, 128)
  AE_TC_CASE(64, 64) AE_TC_CASE(64, 32)
  AE_TC_CASE(128, 256) AE_TC_CASE(256, 128) AE_TC_CASE(128, 64) AE_TC_CASE(256, 64)
#undef AE_TC_CASE
  return AETHER_E_UNSUPPORTED;
}

}  // namespace aether
