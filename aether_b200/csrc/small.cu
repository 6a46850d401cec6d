// Whole inference rollouts of small scenes in ONE launch (BASELINE configs 1-3).
//   step    : nn/state2state/aether.py:169-186 (field, frames, edge features, 4 message-passing layers, out MLP, Euler)
//   caller  : experiments/lorentz/main.py:243-247 (edge_attr = [q_i q_j, |x_i - x_j|]), v_{t+1} = x_{t+1} - x_t
//   kNN     : experiments/ind/single_ind_data.py:186-217, rebuilt from x_t every step
//
// One CTA of 512 threads owns one scene (n <= 64 particles) for all T steps.  Nothing of a step leaves the SM except the
// new positions: the edge state [E, 64], the node state, the frames and the edge list live in shared memory, and
// the kNN lists are rebuilt there from the positions the step just produced.  The multi-launch path (step.cu) needs ~35
// launches per step at these sizes and is launch-latency bound even inside a CUDA graph (170 us per step at 16 x 5
// particles).
//
// Weights.  A step multiplies by 32 weight tiles of 64 x 64 (per layer: W_e, W_2, the two halves of W_3 and of W_4, W_s
// and W_r of the next layer; then the out MLP's two).  A pack launch in front of the rollout writes them once as
// ready-to-use shared-memory images (512 KB fp32 / 256 KB bf16, L2-resident and shared by every CTA); the rollout kernel
// streams them through a ring of 3..16 shared-memory buffers with cp.async.bulk + mbarrier (one thread issues, nobody
// copies), as many tiles ahead as the scene's size leaves room for.
//
// Two arithmetic modes (template parameter BF16, chosen by the model's precision flag):
//   fp32 (tolerance 1e-5): CUDA cores.  A warp owns 2 RW rows of a row chunk -- even rows on lanes 0-15, odd rows on
//     lanes 16-31 -- and each lane four adjacent output columns: per 4 steps of k that is 4 + RW 128-bit shared loads
//     for 16 RW FMAs.  The A operands (edge state, node state) are stored with their 16-byte chunks XOR-swizzled by the
//     row parity, so the two rows a warp reads per instruction sit in different banks.
//   tensor cores (tolerance 2e-2): every GEMM on mma.sync.m16n8k16 (bf16 operands, fp32 accumulate) with ldmatrix on
//     128-byte rows swizzled by row & 7; the edge state is bf16 (twice the edges per scene), 256-edge chunks with 16
//     rows per warp, node GEMMs with warp = (16-row tile, 16-column quarter) on bf16 copies written by the producing
//     phases.  Aggregation, residuals, field network, geometry and the Euler update stay fp32.
// Compiled with -fmad=false like the other geometry files; the GEMM loops use explicit fmaf.
#include <type_traits>

#include "geom.cuh"
#include "param_slots.cuh"
#include "tc_common.cuh"

namespace aether {
namespace {

using tc::smem_u32;

constexpr int SH = 64;            // hidden size this kernel is built for
constexpr int ST = 512;           // threads per CTA
constexpr int SWARPS = ST / 32;
constexpr int CH = 128;           // edge rows per chunk of the chained edge MLP
constexpr int FH = 32, EMB = 16;  // field network
constexpr int kSmallMaxNodes = 64;
constexpr int kSmemLimit = 232448;
constexpr int kTile = SH * SH;    // floats per packed weight tile
constexpr int kTilesPerStep = 32;
constexpr int kMaxRing = 16;

struct SmallParams { const float* p[P_NUM]; };

struct SmallArgs {
  const int32_t* node_off;        // [S+1] first particle of every scene
  const int64_t* edge_off;        // [S+1] first edge of every scene
  const int32_t* order;           // [S] scene handled by CTA i (largest first), or null: identity
  const int64_t* send;            // static edge list (global ids, grouped by scene) or null: kNN
  const int64_t* recv;
  const float* x0; const float* v0; const float* q;
  const float* tiles;             // [32][64][64] packed weight tiles
  float* traj;                    // [T, Nn, D]
  int64_t* send_out; int64_t* recv_out;   // optional: the edge list of the last step (global ids)
  int64_t n_nodes;
  int32_t steps, knn_k, max_nodes, max_edges, ring;
};

// bytes of a packed tile: 64 x 64 fp32, or bf16 in the tensor-core mode
__host__ __device__ inline uint32_t small_tile_bytes(bool bf16) { return bf16 ? kTile * 2 : kTile * 4; }

// shared-memory plan, in floats (all regions 16-byte aligned)
struct SmallPlan {
  int S, T1, ring, X, PS, PR, x, v, xn, f, pred, R, x0, q, send, recv, elist, rowptr, bars, bias, n16, total, nb, slot;
};
__host__ __device__ inline int up4(int v) { return (v + 3) & ~3; }
// bf16 = 1: edge state and chunk temporary as bf16 (128-byte rows: half the floats), T1 then holds 256 rows
__host__ __device__ inline SmallPlan small_plan(int nmax, int emax, int nb, int bf16) {
  SmallPlan p;
  int o = 0;
  auto take = [&](int n) { const int r = (o + 31) & ~31; o = r + up4(n); return r; };     // 128-byte aligned regions
  p.S = take((emax > 0 ? emax : 1) * (bf16 ? SH / 2 : SH));
  p.T1 = take(CH * SH);
  p.X = take(nmax * SH); p.PS = take(nmax * SH); p.PR = take(nmax * SH);
  p.x = take(nmax * 3); p.v = take(nmax * 3); p.xn = take(nmax * 3); p.f = take(nmax * 3); p.pred = take(nmax * 3);
  p.R = take(nmax * 9); p.x0 = take(nmax * 9); p.q = take(nmax);
  p.send = take(emax); p.recv = take(emax); p.elist = take(emax); p.rowptr = take(nmax + 1);
  p.bars = take(2 * kMaxRing);
  p.bias = take(2 * SH);
  p.n16 = take(bf16 ? 4 * nmax * (SH / 2) : 0);      // bf16 A operands of the node GEMMs: y, x, two hidden halves
  o = (o + 31) & ~31;
  p.ring = o;
  p.slot = bf16 ? kTile / 2 : kTile;                  // floats per ring buffer (bf16 tiles are 8 KB)
  if (nb <= 0) {                    // as many ring buffers as fit
    nb = (kSmemLimit / 4 - o) / p.slot;
    if (nb > kMaxRing) nb = kMaxRing;
  }
  p.nb = nb;
  p.total = o + (nb > 0 ? nb : 0) * p.slot;
  return p;
}

struct TileDesc { const float* W; int ldw, K; };
// tile i of a step: layer l = i / 8 + 1 has W_e, W_2, W_3[0:H], W_3[H:2H], W_4[:, 0:H], W_4[:, H:2H] and (l < 4) the next
// layer's W_s, W_r; tiles 30, 31 are the out MLP's first two Linears
__host__ __device__ inline TileDesc small_tile(const SmallParams& P, int i, int FE) {
  if (i >= 30) return {P.p[i == 30 ? P_OUT0_W : P_OUT3_W], SH, SH};
  const int l = i / 8 + 1, j = i - 8 * (l - 1);
  const LayerSlots s = layer_slots(l);
  switch (j) {
    case 0: if (l == 1) return {P.p[s.m0w], FE, FE}; return {P.p[s.m0w] + 2 * SH, 3 * SH, SH};
    case 1: return {P.p[s.m2w], SH, SH};
    case 2: return {P.p[s.u0w], SH, SH};
    case 3: return {P.p[s.u0w] + SH * SH, SH, SH};
    case 4: return {P.p[s.u2w], 2 * SH, SH};
    case 5: return {P.p[s.u2w] + SH, 2 * SH, SH};
    case 6: return {P.p[layer_slots(l + 1).m0w], 3 * SH, SH};
    default: return {P.p[layer_slots(l + 1).m0w] + SH, 3 * SH, SH};
  }
}

// tiles[i][k][n] = W_i[n][k] (zero rows for k >= K): one block per tile, transposed through shared memory
// Tensor-core mode: the tiles are written as bf16 [n][k] rows of 128 bytes with their 16-byte chunks XOR-swizzled by
// n & 7 -- exactly the shared-memory image ldmatrix reads -- and layer 1's W_e as [W | W] against the [hi | lo] features.
__global__ void __launch_bounds__(256) small_pack_kernel(const __grid_constant__ SmallParams P, int FE, float* __restrict__ tiles,
                                                         int bf16) {
  __shared__ float t[SH][SH + 1];
  const TileDesc d = small_tile(P, blockIdx.x, FE);
  if (bf16) {
    __nv_bfloat16* o16 = reinterpret_cast<__nv_bfloat16*>(tiles + (size_t)blockIdx.x * kTile);
    const bool first = blockIdx.x == 0;
    for (int i = threadIdx.x; i < kTile; i += 256) {
      const int n = i >> 6, k = i & 63, ks = first ? (k & 31) : k;
      const float v = ks < d.K ? d.W[(size_t)n * d.ldw + ks] : 0.0f;
      o16[n * 64 + ((((k >> 3) ^ (n & 7)) << 3) | (k & 7))] = __float2bfloat16_rn(v);
    }
    return;
  }
  for (int i = threadIdx.x; i < kTile; i += 256) {
    const int n = i >> 6, k = i & 63;
    t[k][n] = k < d.K ? d.W[(size_t)n * d.ldw + k] : 0.0f;
  }
  __syncthreads();
  float* out = tiles + (size_t)blockIdx.x * kTile;
  for (int i = threadIdx.x; i < kTile; i += 256) out[i] = t[i >> 6][i & 63];
}

// SiLU through the fast exponential and division (ex2.approx / rcp.approx paths, ~2 ulp each): the IEEE forms cost ~35
// instructions per element and were a quarter of all instructions of the kernel
__device__ __forceinline__ float silu_s(float z) { return __fdividef(z, 1.0f + __expf(-z)); }
__device__ __forceinline__ float4 silu4(float4 v) { return make_float4(silu_s(v.x), silu_s(v.y), silu_s(v.z), silu_s(v.w)); }
__device__ __forceinline__ float4 add4(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
// column of element (r, c) in an A-operand array: 16-byte chunks swizzled by the row parity
__device__ __forceinline__ int swz(int r, int c) { return c ^ ((r & 1) << 2); }

// acc[i][:] += sum_k A[r_i][k] * Wt[k][c4 .. c4+3], r_i = r0 + 2 i + (lane >> 4), r0 even; rows past M - 1 are clamped
// (their results are dropped by the caller)
template <int RW>
__device__ __forceinline__ void gemm_acc(float (&acc)[RW][4], const float* __restrict__ A, int lda, int r0, int M, int K,
                                         const float* __restrict__ Wt, int lane) {
  const int gi = lane >> 4;
  int ao[RW];
#pragma unroll
  for (int i = 0; i < RW; ++i) {
    const int r = r0 + 2 * i + gi;
    ao[i] = (r < M ? r : M - 1) * lda;
  }
  const int sx = gi << 2;
  const float* w = Wt + (lane & 15) * 4;
#pragma unroll (RW == 1 ? 4 : 2)
  for (int k = 0; k < K; k += 4) {
    float4 wv[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) wv[q] = *reinterpret_cast<const float4*>(w + (k + q) * SH);
#pragma unroll
    for (int i = 0; i < RW; ++i) {
      const float4 a = *reinterpret_cast<const float4*>(A + ao[i] + (k ^ sx));
      acc[i][0] = fmaf(a.x, wv[0].x, acc[i][0]); acc[i][1] = fmaf(a.x, wv[0].y, acc[i][1]);
      acc[i][2] = fmaf(a.x, wv[0].z, acc[i][2]); acc[i][3] = fmaf(a.x, wv[0].w, acc[i][3]);
      acc[i][0] = fmaf(a.y, wv[1].x, acc[i][0]); acc[i][1] = fmaf(a.y, wv[1].y, acc[i][1]);
      acc[i][2] = fmaf(a.y, wv[1].z, acc[i][2]); acc[i][3] = fmaf(a.y, wv[1].w, acc[i][3]);
      acc[i][0] = fmaf(a.z, wv[2].x, acc[i][0]); acc[i][1] = fmaf(a.z, wv[2].y, acc[i][1]);
      acc[i][2] = fmaf(a.z, wv[2].z, acc[i][2]); acc[i][3] = fmaf(a.z, wv[2].w, acc[i][3]);
      acc[i][0] = fmaf(a.w, wv[3].x, acc[i][0]); acc[i][1] = fmaf(a.w, wv[3].y, acc[i][1]);
      acc[i][2] = fmaf(a.w, wv[3].z, acc[i][2]); acc[i][3] = fmaf(a.w, wv[3].w, acc[i][3]);
    }
  }
}

template <int RW> struct Acc { float v[RW][4]; };
template <int RW>
__device__ __forceinline__ void acc_zero(Acc<RW>& a) {
#pragma unroll
  for (int i = 0; i < RW; ++i) a.v[i][0] = a.v[i][1] = a.v[i][2] = a.v[i][3] = 0.0f;
}
// epi(row, c4, float4) for the rows of [row0, M) this lane owns
template <int RW, class Epi>
__device__ __forceinline__ void acc_store(const Acc<RW>& a, int r0, int M, Epi&& epi) {
  const int lane = threadIdx.x & 31, gi = lane >> 4, c4 = (lane & 15) * 4;
#pragma unroll
  for (int i = 0; i < RW; ++i) {
    const int r = r0 + 2 * i + gi;
    if (r < M) epi(r, c4, make_float4(a.v[i][0], a.v[i][1], a.v[i][2], a.v[i][3]));
  }
}
// one GEMM over rows [row0, M) with 2 RW rows per warp
template <int RW, class Epi>
__device__ __forceinline__ void gemm_rows(const float* A, int lda, int K, const float* Wt, int row0, int M, Epi&& epi) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r0 = row0 + warp * 2 * RW;
  if (r0 >= M) return;
  Acc<RW> a;
  acc_zero(a);
  gemm_acc<RW>(a.v, A, lda, r0, M, K, Wt, lane);
  acc_store<RW>(a, r0, M, epi);
}
// rows [row0, row0 + rows), rows <= 128: the rows-per-warp that covers them with 16 warps
template <class Epi>
__device__ __forceinline__ void chunk_gemm(const float* A, int lda, int K, const float* Wt, int row0, int rows, Epi&& epi) {
  const int M = row0 + rows;
  if (rows <= 2 * SWARPS) gemm_rows<1>(A, lda, K, Wt, row0, M, epi);
  else if (rows <= 4 * SWARPS) gemm_rows<2>(A, lda, K, Wt, row0, M, epi);
  else gemm_rows<4>(A, lda, K, Wt, row0, M, epi);
}

// ---- tensor-core mode (models with precision = "bf16"): the edge MLP on mma.sync m16n8k16, bf16 operands, fp32 accumulate --
// Edge state and chunk temporary are [rows][64] bf16 with 128-byte rows whose 16-byte chunks are XOR-swizzled by row & 7
// (conflict-free ldmatrix); the weight tiles arrive in the same layout from the pack kernel.
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// byte offset of 16-byte chunk c of row r
__device__ __forceinline__ uint32_t sw16(int r, int c) { return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4)); }

// acc[j][:] (n-tile j = columns 8j .. 8j+7) = A[r0 .. r0+15][0 .. 63] W^T; rows past rmax are clamped (results dropped).
// Fragment of lane (g = lane / 4, t = lane % 4): acc[j][0..1] = row r0 + g, columns 8j + 2t, +1; acc[j][2..3] = row r0 + g + 8.
__device__ __forceinline__ void mma_rows16(float (&acc)[8][4], uint32_t aBase, int r0, int rmax, uint32_t wBase, int lane) {
  int ar = r0 + (lane & 7) + ((lane >> 3) & 1) * 8;
  ar = ar < rmax ? ar : rmax;
  const int ac = lane >> 4;
  const int bn = (lane & 7) + ((lane >> 4) << 3), bc = (lane >> 3) & 1;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    uint32_t af[4];
    ldsm_x4(aBase + sw16(ar, 2 * ks + ac), af);
#pragma unroll
    for (int jp = 0; jp < 4; ++jp) {
      uint32_t bf[4];
      ldsm_x4(wBase + sw16(16 * jp + bn, 2 * ks + bc), bf);
      mma_bf16(acc[2 * jp], af, bf[0], bf[1]);
      mma_bf16(acc[2 * jp + 1], af, bf[2], bf[3]);
    }
  }
}

template <int D, bool BF16>
__global__ void __launch_bounds__(ST, 1) small_rollout_kernel(const __grid_constant__ SmallParams P, const SmallArgs a) {
  constexpr int O = D * (D - 1) / 2, FE = 7 * D + O + 2, LD = (FE + 3) / 4 * 4, FI = 2 * D + EMB;
  extern __shared__ __align__(128) float sm[];
  const SmallPlan pl = small_plan(a.max_nodes, a.max_edges, a.ring, BF16 ? 1 : 0);
  float* const S = sm + pl.S;            // [E, 64] edge state: features (layer 1), then silu(pre2_l)      (swizzled rows)
  float* const T1 = sm + pl.T1;          // chunk temporary [128, 64] (swizzled); kNN distances; field hidden layers
  float* const Y = T1;                   // y_l [n, 64] (swizzled): live from the aggregation to the node update, when
                                         // nothing else uses T1
  float* const X = sm + pl.X;            // x_l [n, 64] (swizzled)
  float* const PS = sm + pl.PS;          // sender / receiver terms of the factorised first Linear [n, 64] each
  float* const PR = sm + pl.PR;
  float* const HID = PS;                 // [n, 128] hidden layer of the node update / out MLP temp (swizzled): P_s | P_r
                                         // of a layer are dead once its edge MLP is done
  float* const sx = sm + pl.x;
  float* const sv = sm + pl.v;
  float* const sxn = sm + pl.xn;
  float* const sf = sm + pl.f;
  float* const sPred = sm + pl.pred;
  float* const sR = sm + pl.R;
  float* const sx0 = sm + pl.x0;
  float* const sq = sm + pl.q;
  int* const sSend = reinterpret_cast<int*>(sm + pl.send);
  int* const sRecv = reinterpret_cast<int*>(sm + pl.recv);
  int* const elist = reinterpret_cast<int*>(sm + pl.elist);
  int* const rowptr = reinterpret_cast<int*>(sm + pl.rowptr);
  float* const sBias = sm + pl.bias;
  float* const ringbuf = sm + pl.ring;
  const uint32_t bar0 = smem_u32(sm + pl.bars);
  const int NB = pl.nb;
  constexpr int SLOT = BF16 ? kTile / 2 : kTile;
  // tensor-core mode: bf16 copies of the node GEMMs' A operands, [n][64] with the 128-byte swizzle
  char* const Y16 = reinterpret_cast<char*>(sm + pl.n16);
  char* const X16 = Y16 + a.max_nodes * 128;
  char* const HA16 = X16 + a.max_nodes * 128;
  char* const HB16 = HA16 + a.max_nodes * 128;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int sc = a.order ? a.order[blockIdx.x] : blockIdx.x;
  const int n0 = a.node_off[sc], n = a.node_off[sc + 1] - n0;
  if (n <= 0) return;
  const int64_t e0 = a.edge_off[sc];
  const int E = (int)(a.edge_off[sc + 1] - e0);
  if (n > a.max_nodes || E > a.max_edges || E < 0 || NB < 3) __trap();
  const bool knn = a.send == nullptr;
  const int kk = knn ? (a.knn_k < n - 1 ? a.knn_k : n - 1) : 0;
  if (knn && n * kk != E) __trap();

  // ---- weight-tile ring: tile g (g counts over all steps) lives in buffer g % NB, its landing barrier completes phase
  //      g / NB.  Thread 0 issues; a buffer is re-armed only after a __syncthreads that follows its last reader.
  if (tid == 0) {
    for (int b = 0; b < NB; ++b) tc::mbar_init(bar0 + 8 * b, 1);
    tc::fence_mbar_init();
  }
  const int total_tiles = a.steps * kTilesPerStep;
  int issued = 0;                       // uniform over the CTA
  auto ring_fill = [&](int head) {      // tiles < head are released: issue while a buffer is free
    const int lim = head + NB < total_tiles ? head + NB : total_tiles;
    if (tid == 0)
      for (int g = issued; g < lim; ++g) {
        const int b = g % NB;
        const uint32_t nbytes = small_tile_bytes(BF16);
        tc::mbar_arrive_expect_tx(bar0 + 8 * b, nbytes);
        tc::bulk_load(smem_u32(ringbuf + b * SLOT), a.tiles + (size_t)(g % kTilesPerStep) * kTile, nbytes, bar0 + 8 * b);
      }
    if (lim > issued) issued = lim;
  };
  auto tile = [&](int g) -> const float* { return ringbuf + (g % NB) * SLOT; };
  // wait until tile g has landed, then barrier (which also orders the shared-memory writes of the phase before)
  auto wait_tile = [&](int g) { tc::mbar_wait(bar0 + 8 * (g % NB), (uint32_t)((g / NB) & 1)); };
  auto acquire = [&](int g) {
    wait_tile(g);
    __syncthreads();
  };

  for (int i = tid; i < n * D; i += ST) { sx[i] = a.x0[(size_t)n0 * D + i]; sv[i] = a.v0[(size_t)n0 * D + i]; }
  for (int i = tid; i < n; i += ST) sq[i] = a.q[n0 + i];
  if (!knn)
    for (int e = tid; e < E; e += ST) {
      long long s = a.send[e0 + e] - n0, r = a.recv[e0 + e] - n0;       // clamped for memory safety (host validates)
      s = s < 0 ? 0 : (s >= n ? n - 1 : s);
      r = r < 0 ? 0 : (r >= n ? n - 1 : r);
      sSend[e] = (int)s; sRecv[e] = (int)r;
    }
  __syncthreads();                      // barriers initialised, state loaded
  ring_fill(0);

  // in-edges of every receiver in ascending edge order (= the stable receiver sort of graph.cu): one warp per receiver
  auto build_csr = [&]() {
    for (int r = warp; r < n; r += SWARPS) {
      int cnt = 0;
      for (int eb = 0; eb < E; eb += 32) {
        const int e = eb + lane;
        cnt += __popc(__ballot_sync(0xffffffffu, e < E && sRecv[e] == r));
      }
      if (lane == 0) rowptr[r + 1] = cnt;
    }
    __syncthreads();
    if (tid == 0) {
      rowptr[0] = 0;
      for (int r = 0; r < n; ++r) rowptr[r + 1] += rowptr[r];
    }
    __syncthreads();
    for (int r = warp; r < n; r += SWARPS) {
      int pos = rowptr[r];
      for (int eb = 0; eb < E; eb += 32) {
        const int e = eb + lane;
        const bool m = e < E && sRecv[e] == r;
        const unsigned b = __ballot_sync(0xffffffffu, m);
        if (m) elist[pos + __popc(b & ((1u << lane) - 1u))] = e;
        pos += __popc(b);
      }
    }
    __syncthreads();
  };
  if (!knn) build_csr();

  // tensor-core node GEMMs: warp = (row tile nrt of 16 particles, column quarter ncq of 16 outputs); acc[j] = n-tile j of
  // the quarter.  Lane (g, tq): acc[j][0..1] = row 16 nrt + g, columns 16 ncq + 8 j + 2 tq, +1; acc[j][2..3] = row + 8.
  const int nrt = warp & 3, ncq = warp >> 2;
  const bool nact = nrt * 16 < n;
  auto node_mma = [&](float (&acc)[2][4], uint32_t aBase, uint32_t wBase, bool zero) {
    if (zero) {
#pragma unroll
      for (int j = 0; j < 2; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0f;
    }
    int ar = nrt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
    ar = ar < n - 1 ? ar : n - 1;
    const int ac = lane >> 4, bn = 16 * ncq + (lane & 7) + ((lane >> 4) << 3), bc = (lane >> 3) & 1;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      uint32_t af[4], bf[4];
      ldsm_x4(aBase + sw16(ar, 2 * ks + ac), af);
      ldsm_x4(wBase + sw16(bn, 2 * ks + bc), bf);
      mma_bf16(acc[0], af, bf[0], bf[1]);
      mma_bf16(acc[1], af, bf[2], bf[3]);
    }
  };
  auto node_each = [&](const float (&acc)[2][4], auto&& f) {
    const int g = lane >> 2, tq = lane & 3;
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      const int r = nrt * 16 + g + 8 * hf;
      if (r < n) {
#pragma unroll
        for (int j = 0; j < 2; ++j) f(r, 16 * ncq + 8 * j + 2 * tq, acc[j][2 * hf], acc[j][2 * hf + 1]);
      }
    }
  };

  for (int t = 0; t < a.steps; ++t) {
    const int g0 = t * kTilesPerStep;   // first tile of this step
    if (knn) {
      // ---- k nearest neighbours of every agent from x_t: rank of (d, j) among the agent's candidates -----------------
      for (int idx = tid; idx < n * n; idx += ST) {
        const int i = idx / n, j = idx - i * n;
        T1[idx] = knn_dist(sx[i * D], sx[i * D + 1], sx[j * D], sx[j * D + 1]);
      }
      __syncthreads();
      const bool dump = a.send_out != nullptr && t == a.steps - 1;
      for (int idx = tid; idx < n * n; idx += ST) {
        const int i = idx / n, j = idx - i * n;
        if (i == j) continue;
        const float d = T1[idx];
        int rank = 0;
        for (int jj = 0; jj < n; ++jj) {
          const float dd = T1[i * n + jj];
          rank += (jj != i) && (dd < d || (dd == d && jj < j));
        }
        if (rank < kk) {
          sSend[i * kk + rank] = i; sRecv[i * kk + rank] = j;
          if (dump) { a.send_out[e0 + i * kk + rank] = n0 + i; a.recv_out[e0 + i * kk + rank] = n0 + j; }
        }
      }
      __syncthreads();
      build_csr();
    }

    // ---- neural field at the particles (aether.py:108-134).  Its weights (7.6 KB) go through the free upper half of T1,
    //      first Linears transposed, so that the lanes (one output unit each) read consecutive words -----------------------
    {
      float* a1 = T1;
      float* a2 = T1 + kSmallMaxNodes * FH;
      float* W0t = T1 + 2 * kSmallMaxNodes * FH;        // [FI][32]
      float* W2t = W0t + FI * FH;                       // [32][32]
      float* W4s = W2t + FH * FH;                       // [D][32]
      float* fb = W4s + D * FH;                         // b0 | b2 | b4
      float* emb = fb + 2 * FH + 4;                     // [3][16]
      for (int i = tid; i < FH * FI; i += ST) { const int o = i / FI, k = i - o * FI; W0t[k * FH + o] = __ldg(P.p[P_F0_W] + i); }
      for (int i = tid; i < FH * FH; i += ST) W2t[(i & 31) * FH + (i >> 5)] = __ldg(P.p[P_F2_W] + i);
      for (int i = tid; i < D * FH; i += ST) W4s[i] = __ldg(P.p[P_F4_W] + i);
      if (tid < FH) { fb[tid] = __ldg(P.p[P_F0_B] + tid); fb[FH + tid] = __ldg(P.p[P_F2_B] + tid); }
      if (tid < D) fb[2 * FH + tid] = __ldg(P.p[P_F4_B] + tid);
      if (tid < 3 * EMB) emb[tid] = __ldg(P.p[P_EMB] + tid);
      __syncthreads();
      for (int idx = tid; idx < n * FH; idx += ST) {
        const int node = idx >> 5, o = idx & 31;
        long long ci = (long long)(sq[node] + 1.0f);
        ci = ci < 0 ? 0 : (ci > 2 ? 2 : ci);
        const float* em = emb + ci * EMB;
        float s = fb[o];
#pragma unroll
        for (int k = 0; k < D; ++k) s = fmaf(sx[node * D + k], W0t[k * FH + o], s);
#pragma unroll
        for (int k = 0; k < D; ++k) s = fmaf(sv[node * D + k], W0t[(D + k) * FH + o], s);
#pragma unroll
        for (int k = 0; k < EMB; ++k) s = fmaf(em[k], W0t[(2 * D + k) * FH + o], s);
        a1[idx] = silu_s(s);
      }
      __syncthreads();
      for (int idx = tid; idx < n * FH; idx += ST) {
        const int node = idx >> 5, o = idx & 31;
        float s = fb[FH + o];
#pragma unroll
        for (int k = 0; k < FH; ++k) s = fmaf(a1[node * FH + k], W2t[k * FH + o], s);
        a2[idx] = silu_s(s);
      }
      __syncthreads();
      for (int idx = tid; idx < n * D; idx += ST) {
        const int node = idx / D, d = idx - node * D;
        float s = fb[2 * FH + d];
#pragma unroll
        for (int k = 0; k < FH; ++k) s = fmaf(a2[node * FH + k], W4s[d * FH + k], s);
        sf[idx] = s;
      }
      __syncthreads();
    }
    // ---- frames, node features (aether.py:33-50) ---------------------------------------------------------------------
    for (int node = tid; node < n; node += ST) node_frame_body<D>(node, sv, sf, sR, sx0);
    __syncthreads();
    // ---- caller's edge attributes (main.py:243-246) and the local-frame edge features (aether.py:81-99,177) -----------
    for (int e = tid; e < E; e += ST) {
      const int j = sSend[e], i = sRecv[e];
      float ss = 0.0f;
#pragma unroll
      for (int k = 0; k < D; ++k) { const float dd = sx[j * D + k] - sx[i * D + k]; ss = ss + dd * dd; }
      float out[32];
      edge_features_core<D>(j, i, sx, sv, sf, sR, sx0, sq[j] * sq[i], sqrtf(ss), out);
#pragma unroll
      for (int k = LD; k < 32; ++k) out[k] = 0.0f;
      if constexpr (BF16) {
        // [hi(32) | lo(32)] bf16, hi = bf16(v), lo = bf16(v - hi): ~16 bits of every feature reach the GEMM (geom.cu)
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const __nv_bfloat16 h0 = __float2bfloat16_rn(out[2 * k]), h1 = __float2bfloat16_rn(out[2 * k + 1]);
          const __nv_bfloat16 l0 = __float2bfloat16_rn(out[2 * k] - __bfloat162float(h0));
          const __nv_bfloat16 l1 = __float2bfloat16_rn(out[2 * k + 1] - __bfloat162float(h1));
          hi[k] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
          lo[k] = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
        }
        char* row = reinterpret_cast<char*>(S);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          *reinterpret_cast<uint4*>(row + sw16(e, k)) = make_uint4(hi[4 * k], hi[4 * k + 1], hi[4 * k + 2], hi[4 * k + 3]);
          *reinterpret_cast<uint4*>(row + sw16(e, 4 + k)) = make_uint4(lo[4 * k], lo[4 * k + 1], lo[4 * k + 2], lo[4 * k + 3]);
        }
      } else {
      float4* dst = reinterpret_cast<float4*>(S + e * SH);
#pragma unroll
      for (int k = 0; k < 8; ++k) dst[k ^ (e & 1)] = make_float4(out[4 * k], out[4 * k + 1], out[4 * k + 2], out[4 * k + 3]);
      }
    }
    // (the barrier of the first acquire below orders these writes)

    // ---- four message-passing layers (locs.py:187-190, 227-243) ------------------------------------------------------
    const int c4 = (lane & 15) * 4;
    auto ld4 = [&](const float* p) { return __ldg(reinterpret_cast<const float4*>(p + c4)); };
    for (int l = 1; l <= 4; ++l) {
      const LayerSlots sl = layer_slots(l);
      const int b = g0 + 8 * (l - 1);   // tiles b .. b+7 (b+5 for the last layer)
      if constexpr (BF16) {
        // tensor-core mode: 256-edge chunks, 16 rows per warp, both GEMMs on mma.sync; biases through shared memory
        if (tid < SH) { sBias[tid] = __ldg(P.p[sl.m0b] + tid); sBias[SH + tid] = __ldg(P.p[sl.m2b] + tid); }
        wait_tile(b);
        acquire(b + 1);
        if (b > 0) ring_fill(b);
        const uint32_t we = smem_u32(tile(b)), w2 = smem_u32(tile(b + 1));
        const uint32_t s16 = smem_u32(S), t16 = smem_u32(T1);
        char* const Sb = reinterpret_cast<char*>(S);
        char* const Tb = reinterpret_cast<char*>(T1);
        const int g = lane >> 2, tq = lane & 3;
        for (int row0 = 0; row0 < E; row0 += 2 * CH) {
          const int rows = E - row0 < 2 * CH ? E - row0 : 2 * CH;
          const int lr0 = 16 * warp, r0 = row0 + lr0;
          float acc[8][4];
          if (lr0 < rows) {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0f;
            mma_rows16(acc, s16, r0, E - 1, we, lane);
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
              const int lr = lr0 + g + 8 * hf, r = row0 + lr;
              if (r < E) {
                const float* ps = PS + sSend[r] * SH + 2 * tq;
                const float* pr = PR + sRecv[r] * SH + 2 * tq;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  float v0 = acc[j][2 * hf], v1 = acc[j][2 * hf + 1];
                  if (l == 1) { v0 += sBias[8 * j + 2 * tq]; v1 += sBias[8 * j + 2 * tq + 1]; }
                  else {
                    const float2 a2 = *reinterpret_cast<const float2*>(ps + 8 * j);
                    const float2 c2 = *reinterpret_cast<const float2*>(pr + 8 * j);
                    v0 += a2.x + c2.x; v1 += a2.y + c2.y;
                  }
                  *reinterpret_cast<uint32_t*>(Tb + sw16(lr, j) + 4 * tq) = tc::pack_bf16(tc::silu_fast(v0), tc::silu_fast(v1));
                }
              }
            }
          }
          __syncthreads();
          if (lr0 < rows) {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0f;
            mma_rows16(acc, t16, lr0, rows - 1, w2, lane);
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
              const int r = r0 + g + 8 * hf;
              if (r < E) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float v0 = acc[j][2 * hf] + sBias[SH + 8 * j + 2 * tq], v1 = acc[j][2 * hf + 1] + sBias[SH + 8 * j + 2 * tq + 1];
                  *reinterpret_cast<uint32_t*>(Sb + sw16(r, j) + 4 * tq) = tc::pack_bf16(tc::silu_fast(v0), tc::silu_fast(v1));
                }
              }
            }
          }
          __syncthreads();
        }
        ring_fill(b + 2);
      } else
      {
        const float4 b1 = ld4(P.p[sl.m0b]), b2 = ld4(P.p[sl.m2b]);
        wait_tile(b);
        acquire(b + 1);                 // W_e and W_2; the barrier also orders the features / P_s, P_r writes
        if (b > 0) ring_fill(b);        // the previous layer's (or step's) last tile is released by that barrier
        const float* We = tile(b);
        const float* W2 = tile(b + 1);
        // edge MLP, chained per chunk of 128 edges: pre1 -> SiLU -> pre2 -> SiLU, in place in the edge state
        for (int row0 = 0; row0 < E; row0 += CH) {
          const int rows = E - row0 < CH ? E - row0 : CH;
          if (l == 1)
            chunk_gemm(S, SH, 32, We, row0, rows, [&](int r, int c, float4 v) {
              *reinterpret_cast<float4*>(T1 + (r - row0) * SH + swz(r, c)) = silu4(add4(v, b1));
            });
          else       // b1 is inside P_r (added by the node GEMM that produced it)
            chunk_gemm(S, SH, SH, We, row0, rows, [&](int r, int c, float4 v) {
              const float4 ps = *reinterpret_cast<const float4*>(PS + sSend[r] * SH + c);
              const float4 pr = *reinterpret_cast<const float4*>(PR + sRecv[r] * SH + c);
              *reinterpret_cast<float4*>(T1 + (r - row0) * SH + swz(r, c)) = silu4(add4(v, add4(ps, pr)));
            });
          __syncthreads();
          chunk_gemm(T1, SH, SH, W2, 0, rows, [&](int r, int c, float4 v) {
            *reinterpret_cast<float4*>(S + (row0 + r) * SH + swz(r, c)) = silu4(add4(v, b2));
          });
          __syncthreads();
        }
        ring_fill(b + 2);               // W_e, W_2 released (E = 0: nobody read them)
      }
      // y = res(x) + mean over the in-edges (locs.py:236-240); layer 1: res = Linear(3D -> H) of the node features
      for (int idx = tid; idx < n * SH; idx += ST) {
        const int node = idx >> 6, c = idx & 63;
        const int pb = rowptr[node], pe = rowptr[node + 1];
        float s = 0.0f;
        if constexpr (BF16) {
          const char* Sb = reinterpret_cast<const char*>(S);
          for (int p = pb; p < pe; ++p) {
            const int e = elist[p];
            s += __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(Sb + sw16(e, c >> 3) + 2 * (c & 7)));
          }
        } else {
        for (int p = pb; p < pe; ++p) { const int e = elist[p]; s += S[e * SH + swz(e, c)]; }
        }
        s = s / (float)(pe - pb < 1 ? 1 : pe - pb);
        float base;
        if (l == 1) {
          base = __ldg(P.p[P_L1_RES_B] + c);
          const float* w = P.p[P_L1_RES_W] + c * 3 * D;
#pragma unroll
          for (int k = 0; k < 3 * D; ++k) base = fmaf(sx0[node * 3 * D + k], __ldg(w + k), base);
        } else base = X[node * SH + swz(node, c)];
        Y[node * SH + swz(node, c)] = base + s;
        if constexpr (BF16) *reinterpret_cast<__nv_bfloat16*>(Y16 + sw16(node, c >> 3) + 2 * (c & 7)) = __float2bfloat16_rn(base + s);
      }
      if constexpr (BF16) {
        // tensor-core mode: x = y + W4 silu(W3 y + b3) + b4 and the next layer's node terms on mma.sync; warp = (16-row
        // tile, 16-column quarter), A operands are the bf16 copies written by the producing phases
        const float* b3 = P.p[sl.u0b];
        const float* b4 = P.p[sl.u2b];
        float acc[2][4];
        acquire(b + 2);                 // orders the y writes
        if (nact) {
          node_mma(acc, smem_u32(Y16), smem_u32(tile(b + 2)), true);
          node_each(acc, [&](int r, int c, float v0, float v1) {
            const float2 bb = __ldg(reinterpret_cast<const float2*>(b3 + c));
            *reinterpret_cast<uint32_t*>(HA16 + sw16(r, c >> 3) + 2 * (c & 7)) = tc::pack_bf16(silu_s(v0 + bb.x), silu_s(v1 + bb.y));
          });
        }
        acquire(b + 3);
        ring_fill(b + 3);
        if (nact) {
          node_mma(acc, smem_u32(Y16), smem_u32(tile(b + 3)), true);
          node_each(acc, [&](int r, int c, float v0, float v1) {
            const float2 bb = __ldg(reinterpret_cast<const float2*>(b3 + SH + c));
            *reinterpret_cast<uint32_t*>(HB16 + sw16(r, c >> 3) + 2 * (c & 7)) = tc::pack_bf16(silu_s(v0 + bb.x), silu_s(v1 + bb.y));
          });
        }
        acquire(b + 4);                 // orders the hidden-layer writes
        ring_fill(b + 4);
        if (nact) node_mma(acc, smem_u32(HA16), smem_u32(tile(b + 4)), true);
        acquire(b + 5);
        ring_fill(b + 5);
        if (nact) {
          node_mma(acc, smem_u32(HB16), smem_u32(tile(b + 5)), false);
          node_each(acc, [&](int r, int c, float v0, float v1) {
            const float2 bb = __ldg(reinterpret_cast<const float2*>(b4 + c));
            const int i = r * SH + swz(r, c);
            const float2 y = *reinterpret_cast<const float2*>(Y + i);
            const float x0v = y.x + (v0 + bb.x), x1v = y.y + (v1 + bb.y);
            *reinterpret_cast<float2*>(X + i) = make_float2(x0v, x1v);
            *reinterpret_cast<uint32_t*>(X16 + sw16(r, c >> 3) + 2 * (c & 7)) = tc::pack_bf16(x0v, x1v);
          });
        }
        if (l < 4) {
          const float* bn = P.p[layer_slots(l + 1).m0b];
          acquire(b + 6);               // orders the x writes
          ring_fill(b + 6);
          if (nact) {
            node_mma(acc, smem_u32(X16), smem_u32(tile(b + 6)), true);
            node_each(acc, [&](int r, int c, float v0, float v1) { *reinterpret_cast<float2*>(PS + r * SH + c) = make_float2(v0, v1); });
          }
          acquire(b + 7);
          ring_fill(b + 7);
          if (nact) {
            node_mma(acc, smem_u32(X16), smem_u32(tile(b + 7)), true);
            node_each(acc, [&](int r, int c, float v0, float v1) {
              const float2 bb = __ldg(reinterpret_cast<const float2*>(bn + c));
              *reinterpret_cast<float2*>(PR + r * SH + c) = make_float2(v0 + bb.x, v1 + bb.y);
            });
          }
        }
      } else {
      // x = y + W4 silu(W3 y + b3) + b4 (locs.py:241): hidden layer of 2H in two column halves, one weight tile each
      {
        const float4 b3a = ld4(P.p[sl.u0b]), b3b = ld4(P.p[sl.u0b] + SH), b4 = ld4(P.p[sl.u2b]);
        acquire(b + 2);                 // orders the y writes
        chunk_gemm(Y, SH, SH, tile(b + 2), 0, n, [&](int r, int c, float4 v) {
          *reinterpret_cast<float4*>(HID + r * 2 * SH + swz(r, c)) = silu4(add4(v, b3a));
        });
        acquire(b + 3);
        ring_fill(b + 3);
        chunk_gemm(Y, SH, SH, tile(b + 3), 0, n, [&](int r, int c, float4 v) {
          *reinterpret_cast<float4*>(HID + r * 2 * SH + SH + swz(r, c)) = silu4(add4(v, b3b));
        });
        acquire(b + 4);                 // orders the hidden-layer writes
        ring_fill(b + 4);
        // two K halves accumulated in registers, the second tile awaited in between
        auto u2 = [&](auto rw) {
          constexpr int RW = decltype(rw)::value;
          const int r0 = warp * 2 * RW;
          Acc<RW> acc;
          acc_zero(acc);
          if (r0 < n) gemm_acc<RW>(acc.v, HID, 2 * SH, r0, n, SH, tile(b + 4), lane);
          acquire(b + 5);
          ring_fill(b + 5);
          if (r0 < n) {
            gemm_acc<RW>(acc.v, HID + SH, 2 * SH, r0, n, SH, tile(b + 5), lane);
            acc_store<RW>(acc, r0, n, [&](int r, int c, float4 v) {
              const int i = r * SH + swz(r, c);
              *reinterpret_cast<float4*>(X + i) = add4(*reinterpret_cast<const float4*>(Y + i), add4(v, b4));
            });
          }
        };
        if (n <= 2 * SWARPS) u2(std::integral_constant<int, 1>{}); else u2(std::integral_constant<int, 2>{});
      }
      if (l < 4) {
        // node terms of the next layer's factorised first Linear: P_s = x Ws^T, P_r = x Wr^T + b1
        const float4 bn = ld4(P.p[layer_slots(l + 1).m0b]);
        acquire(b + 6);                 // orders the x writes; the hidden layer (= P_s | P_r storage) is dead
        ring_fill(b + 6);
        chunk_gemm(X, SH, SH, tile(b + 6), 0, n, [&](int r, int c, float4 v) { *reinterpret_cast<float4*>(PS + r * SH + c) = v; });
        acquire(b + 7);
        ring_fill(b + 7);
        chunk_gemm(X, SH, SH, tile(b + 7), 0, n, [&](int r, int c, float4 v) { *reinterpret_cast<float4*>(PR + r * SH + c) = add4(v, bn); });
      }
      }
    }

    // ---- out MLP (locs.py:160-168,193), rotation back and Euler update (local_to_global.py:12-13, aether.py:185) -------
    if constexpr (BF16) {
      float acc[2][4];
      acquire(g0 + 30);                 // orders the x writes
      ring_fill(g0 + 30);
      if (nact) {
        node_mma(acc, smem_u32(X16), smem_u32(tile(g0 + 30)), true);
        node_each(acc, [&](int r, int c, float v0, float v1) {
          const float2 bb = __ldg(reinterpret_cast<const float2*>(P.p[P_OUT0_B] + c));
          *reinterpret_cast<uint32_t*>(HA16 + sw16(r, c >> 3) + 2 * (c & 7)) = tc::pack_bf16(silu_s(v0 + bb.x), silu_s(v1 + bb.y));
        });
      }
      acquire(g0 + 31);
      ring_fill(g0 + 31);
      if (nact) {
        node_mma(acc, smem_u32(HA16), smem_u32(tile(g0 + 31)), true);
        node_each(acc, [&](int r, int c, float v0, float v1) {
          const float2 bb = __ldg(reinterpret_cast<const float2*>(P.p[P_OUT3_B] + c));
          *reinterpret_cast<float2*>(Y + r * SH + swz(r, c)) = make_float2(silu_s(v0 + bb.x), silu_s(v1 + bb.y));
        });
      }
      __syncthreads();
      ring_fill(g0 + 32);
    } else
    {
      const float4 c0 = ld4(P.p[P_OUT0_B]), c3 = ld4(P.p[P_OUT3_B]);
      acquire(g0 + 30);                 // orders the x writes
      ring_fill(g0 + 30);
      chunk_gemm(X, SH, SH, tile(g0 + 30), 0, n, [&](int r, int c, float4 v) {
        *reinterpret_cast<float4*>(HID + r * SH + swz(r, c)) = silu4(add4(v, c0));
      });
      acquire(g0 + 31);
      ring_fill(g0 + 31);
      chunk_gemm(HID, SH, SH, tile(g0 + 31), 0, n, [&](int r, int c, float4 v) {
        *reinterpret_cast<float4*>(Y + r * SH + swz(r, c)) = silu4(add4(v, c3));
      });
      __syncthreads();
      ring_fill(g0 + 32);
    }
    for (int idx = tid; idx < n * D; idx += ST) {
      const int node = idx / D, d = idx - node * D;
      float s = __ldg(P.p[P_OUT6_B] + d);
      const float* w = P.p[P_OUT6_W] + d * SH;
      const int sw = (node & 1) << 2;
#pragma unroll 8
      for (int k = 0; k < SH; ++k) s = fmaf(Y[node * SH + (k ^ sw)], __ldg(w + k), s);
      sPred[idx] = s;
    }
    __syncthreads();
    for (int node = tid; node < n; node += ST) globalize_body<D>(node, sx, sR, sPred, sxn);
    __syncthreads();
    float* out = a.traj + ((size_t)t * a.n_nodes + n0) * D;
    for (int idx = tid; idx < n * D; idx += ST) {
      const float xn = sxn[idx];
      out[idx] = xn;
      sv[idx] = xn - sx[idx];
      sx[idx] = xn;
    }
    __syncthreads();
  }
}

}  // namespace
}  // namespace aether

using namespace aether;

extern "C" size_t aether_small_rollout_scratch_bytes(void) { return (size_t)kTilesPerStep * kTile * sizeof(float); }

extern "C" int aether_small_rollout_supported(int32_t num_dims, int32_t hidden, int32_t max_nodes, int32_t max_edges,
                                              uint32_t flags) {
  if ((num_dims != 2 && num_dims != 3) || hidden != SH || max_nodes < 1 || max_nodes > kSmallMaxNodes || max_edges < 0 ||
      max_edges > (1 << 16))
    return 0;
  return small_plan(max_nodes, max_edges, 0, (flags & AETHER_F_BF16_EDGE_MLP) ? 1 : 0).nb >= 3 ? 1 : 0;
}

template <int D, bool BF16>
static int launch_small(const SmallParams& P, const SmallArgs& a, int n_scenes, size_t smem, cudaStream_t st) {
  AE_CHECK((int)cudaFuncSetAttribute(small_rollout_kernel<D, BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  small_rollout_kernel<D, BF16><<<n_scenes, ST, smem, st>>>(P, a);
  AE_LAUNCH_CHECK();
  return 0;
}

extern "C" int aether_small_rollout(const AetherDims* dims, const float* const* params, int32_t n_scenes,
                                    const int32_t* node_off, const int64_t* edge_off, const int32_t* order,
                                    const int64_t* send, const int64_t* recv, int32_t knn_k, int32_t max_nodes,
                                    int32_t max_edges, const float* x0, const float* v0, const float* charges, int32_t steps,
                                    float* traj, int64_t* send_out, int64_t* recv_out, void* scratch, size_t scratch_bytes,
                                    uint32_t flags, void* stream) {
  if (!dims || !params || !node_off || !edge_off || !x0 || !v0 || !charges || !traj || n_scenes < 0 || steps < 0 ||
      (send == nullptr) != (recv == nullptr) || (send == nullptr && knn_k < 0) || (send_out == nullptr) != (recv_out == nullptr))
    return AETHER_E_BADARG;
  if (!aether_small_rollout_supported(dims->num_dims, dims->hidden, max_nodes, max_edges, flags)) return AETHER_E_UNSUPPORTED;
  if (!scratch || ((uintptr_t)scratch & 127) || scratch_bytes < aether_small_rollout_scratch_bytes()) return AETHER_E_WORKSPACE;
  SmallParams P;
  for (int i = 0; i < P_NUM; ++i) {
    if (!params[i]) return AETHER_E_BADARG;
    if ((uintptr_t)params[i] & 15) return AETHER_E_UNSUPPORTED;     // biases are read as float4
    P.p[i] = params[i];
  }
  if (n_scenes == 0 || steps == 0) return 0;
  const int D = dims->num_dims;
  const bool bf16 = flags & AETHER_F_BF16_EDGE_MLP;
  cudaStream_t st = (cudaStream_t)stream;
  small_pack_kernel<<<kTilesPerStep, 256, 0, st>>>(P, num_edge_feat(D), (float*)scratch, bf16 ? 1 : 0);
  AE_LAUNCH_CHECK();
  const SmallPlan pl = small_plan(max_nodes, max_edges, 0, bf16 ? 1 : 0);
  SmallArgs a;
  a.node_off = node_off; a.edge_off = edge_off; a.order = order; a.send = send; a.recv = recv; a.x0 = x0; a.v0 = v0;
  a.q = charges; a.tiles = (const float*)scratch; a.traj = traj; a.send_out = send_out; a.recv_out = recv_out;
  a.n_nodes = dims->n_nodes; a.steps = steps; a.knn_k = knn_k; a.max_nodes = max_nodes; a.max_edges = max_edges; a.ring = pl.nb;
  const size_t smem = (size_t)pl.total * sizeof(float);
  if (D == 2) return bf16 ? launch_small<2, true>(P, a, n_scenes, smem, st) : launch_small<2, false>(P, a, n_scenes, smem, st);
  return bf16 ? launch_small<3, true>(P, a, n_scenes, smem, st) : launch_small<3, false>(P, a, n_scenes, smem, st);
}
