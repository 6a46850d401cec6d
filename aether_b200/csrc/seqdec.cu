// seq2seq decoder path: RecurrentDecoder.forward (nn/seq2seq/aether.py:589-654), AugmentedLocalizer
// (nn/utils/augmented_global_to_local.py:51-68, nn/utils/canonicalization.py:33-56,111-172) and the field predictor
// (nn/seq2seq/aether.py:83-87).  fp32, inference.  Built with -fmad=false so the geometry rounds like the reference's
// eager ops; the GEMM inner loops ask for FMAs explicitly.
//
// One decoder step is 10 launches (14 with the field predictor), all on the caller's stream, capturable:
//   geometry -> batch A {node terms of msg_fc1 per type, present_msg_fc1 per type} -> tanh of the pair sums
//   -> batch B {msg_fc2 summed over types, present_msg_fc2 summed over types} -> receiver means -> batch C {gate
//   pre-activations} -> gates + state update -> out_mlp.0 -> out_mlp.3 -> out_mlp.6 + Globalizer + inputs.
// msg_fc1 is node-factorised: W1 [recv | send] halves applied to the N node states of a trajectory instead of its
// N(N-1) pairs.  Every "batch" is ONE launch of a multi-problem, multi-term GEMM:
//   C[m,n] = sum over flushed groups of  scale[m] * act( sum_terms A_t[m,:] . W_t[n,:] + bias_t[n] )
// which covers the per-type sums (":608-613, :625-632") and the three-input gate pre-activations (":638-645").
#include "common.cuh"

namespace aether {
namespace {

// Programmatic dependent launch: every kernel of this file is launched with the programmatic-stream-serialization
// attribute, lets its dependents be scheduled when its CTAs are done with their main loops (pdl_trigger; triggering at
// CTA start measured slower: the waiting dependents take SM slots from the primary's later waves) and reads nothing a
// predecessor wrote before pdl_wait returns (= the predecessor grid has completed and its writes are visible).  The
// next kernel's CTAs are then resident and past their set-up when the previous one drains, which takes the launch
// latency out of the 14-link chain of a decoder step (also inside a captured CUDA graph: programmatic edges).
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
cudaError_t seq_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

constexpr float PI_F = 3.14159274f;       // float(np.pi)
constexpr float TWO_PI_F = 6.28318548f;   // float(2 * np.pi)

// ------------------------------------------------------------------------------------------------
// geometry
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void rot3(float th, float ph, float* R) {   // nn/utils/geometry.py:24-34
  const float ct = cosf(th), st = sinf(th), cp = cosf(ph), sp = sinf(ph);
  R[0] = cp * ct; R[1] = -st; R[2] = sp * ct;
  R[3] = cp * st; R[4] = ct;  R[5] = sp * st;
  R[6] = -sp;     R[7] = 0.f; R[8] = cp;
}
__device__ __forceinline__ void spherical3(const float* x, float& rho, float& th, float& ph) {   // geometry.py:37-67
  rho = sqrtf(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
  th = atan2f(x[1], x[0]);
  if (th < 0.f) th = th + TWO_PI_F;
  float c = x[2] / (rho + 1e-7f);
  c = fminf(fmaxf(c, -1.0f), 1.0f);
  ph = acosf(c);
}
// out = R^T v for a row-major 3x3 R
__device__ __forceinline__ void mv3t(const float* R, const float* v, float* o) {
  for (int a = 0; a < 3; ++a) o[a] = R[a] * v[0] + R[3 + a] * v[1] + R[6 + a] * v[2];
}

template <int D> struct SeqGeo {
  static constexpr int O = D * (D - 1) / 2, F = 4 * D + O, NS = 3 * D;
};

// canonicalize_augmented_inputs (canonicalization.py:33-56): x = [pos | vel | field]
template <int D>
__device__ __forceinline__ void canonical(const float* x, float* canon, float* Rinv) {
  if (D == 2) {
    const float a = atan2f(x[3], x[2]);
    const float c = cosf(a), s = sinf(a);
    Rinv[0] = c; Rinv[1] = -s; Rinv[2] = s; Rinv[3] = c;
    canon[0] = 0.f; canon[1] = 0.f;
    canon[2] = sqrtf(x[2] * x[2] + x[3] * x[3]);
    canon[3] = 0.f;
    canon[4] = c * x[4] + s * x[5];
    canon[5] = (-s) * x[4] + c * x[5];
  } else {
    float rho, th, ph;
    spherical3(x + 3, rho, th, ph);
    rot3(th, ph, Rinv);
    canon[0] = canon[1] = canon[2] = 0.f;
    mv3t(Rinv, x + 3, canon + 3);
    mv3t(Rinv, x + 6, canon + 6);
  }
}

// create_augmented_edge_attr_pos_vel / create_augmented_3d_edge_attr_pos_vel (canonicalization.py:111-172):
// sender xj seen from receiver xi
template <int D>
__device__ __forceinline__ void pair_feat(const float* xj, const float* xi, float* o) {
  if (D == 2) {
    const float yaw = atan2f(xi[3], xi[2]);
    const float c = cosf(yaw), s = sinf(yaw);
    const float dx = xj[0] - xi[0], dy = xj[1] - xi[1];
    o[0] = c * dx + s * dy;
    o[1] = (-s) * dx + c * dy;
    float d = atan2f(xj[3], xj[2]) - yaw;                 // angle_diff, geometry.py:117-128
    if (d >= PI_F) d -= TWO_PI_F;
    if (d < -PI_F) d += TWO_PI_F;
    o[2] = d / PI_F;
    o[3] = sqrtf(dx * dx + dy * dy);
    float t = atan2f(dy, dx) - yaw;                       // wrap_angles, geometry.py:108-113
    t = t + (t <= -PI_F ? TWO_PI_F : 0.f);
    t = t - (t > PI_F ? TWO_PI_F : 0.f);
    o[4] = t / PI_F;
    o[5] = c * xj[2] + s * xj[3];
    o[6] = (-s) * xj[2] + c * xj[3];
    o[7] = c * xj[4] + s * xj[5];
    o[8] = (-s) * xj[4] + c * xj[5];
  } else {
    float rho, sy, sp, ry, rp;
    spherical3(xj + 3, rho, sy, sp);
    spherical3(xi + 3, rho, ry, rp);
    float Ri[9], Rj[9];
    rot3(ry, rp, Ri);
    rot3(sy, sp, Rj);
    float dp[3] = {xj[0] - xi[0], xj[1] - xi[1], xj[2] - xi[2]};
    // M = Ri^T Rj^T (":153-154": both factors are the transposed matrices)
    auto M = [&](int a, int b) { return Ri[a] * Rj[3 * b] + Ri[3 + a] * Rj[3 * b + 1] + Ri[6 + a] * Rj[3 * b + 2]; };
    mv3t(Ri, dp, o);
    o[3] = atan2f(M(1, 0), M(0, 0));
    o[4] = asinf(-M(2, 0));
    o[5] = atan2f(M(2, 1), M(2, 2));
    o[6] = sqrtf(dp[0] * dp[0] + dp[1] * dp[1] + dp[2] * dp[2]);
    float r2, th, ph;
    spherical3(o, r2, th, ph);
    o[7] = th;
    o[8] = ph;
    mv3t(Ri, xj + 3, o + 9);
    mv3t(Ri, xj + 6, o + 12);
  }
}

template <int D>
__device__ __forceinline__ void load_state(const float* states, const float* field, int64_t node, float* x) {
  for (int i = 0; i < 2 * D; ++i) x[i] = states[node * 2 * D + i];
  for (int i = 0; i < D; ++i) x[2 * D + i] = field[node * D + i];
}

// Thread per (trajectory, pair) and per (trajectory, particle): edge_attr[B*E, 2F+3D] = [pair | canon(recv) | origin
// pair of recv], rel_feat[B*N, 3D+F] = [canon | origin pair], Rinv[B*N, D*D].
template <int D>
__global__ void seq_geom_kernel(const float* __restrict__ states, const float* __restrict__ field, int N, int64_t B,
                                float* __restrict__ relx, float* __restrict__ Rinv, float* __restrict__ eattr) {
  pdl_wait();
  constexpr int F = SeqGeo<D>::F, NS = SeqGeo<D>::NS;
  const int E = N * (N - 1);
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * (E + N)) return;
  const int64_t b = idx / (E + N);
  const int e = (int)(idx - b * (E + N));
  float origin[NS];
  for (int i = 0; i < NS; ++i) origin[i] = 0.f;
  origin[D] = 1.0f;                                       // augmented_global_to_local.py:41
  float xi[NS], canon[NS], R[D * D], of[F];
  if (e < E) {
    const int send = e / (N - 1), q = e - send * (N - 1), recv = q + (q >= send ? 1 : 0);
    float xj[NS], pf[F];
    load_state<D>(states, field, b * N + recv, xi);
    load_state<D>(states, field, b * N + send, xj);
    pair_feat<D>(xj, xi, pf);
    canonical<D>(xi, canon, R);
    pair_feat<D>(origin, xi, of);
    float* row = eattr + (b * E + e) * (2 * F + NS);
    for (int i = 0; i < F; ++i) row[i] = pf[i];
    for (int i = 0; i < NS; ++i) row[F + i] = canon[i];
    for (int i = 0; i < F; ++i) row[F + NS + i] = of[i];
  } else {
    const int64_t node = b * N + (e - E);
    load_state<D>(states, field, node, xi);
    canonical<D>(xi, canon, R);
    pair_feat<D>(origin, xi, of);
    float* row = relx + node * (NS + F);
    for (int i = 0; i < NS; ++i) row[i] = canon[i];
    for (int i = 0; i < F; ++i) row[NS + i] = of[i];
    for (int i = 0; i < D * D; ++i) Rinv[node * D * D + i] = R[i];
  }
}

// ------------------------------------------------------------------------------------------------
// multi-problem, multi-term fp32 GEMM
// ------------------------------------------------------------------------------------------------
constexpr int GT = 256;
constexpr int MAX_TERMS = 4, MAX_PROB = 9;     // sizeof(Batch) stays under the 4 KB of kernel parameters
enum SeqAct { ACT_NONE = 0, ACT_TANH = 1, ACT_RELU = 2, ACT_SIGMOID = 3, ACT_SILU = 4, ACT_ELU = 5 };

struct Term {
  const float* A;       // [M, K], leading dimension lda
  const float* W;       // [N, K], leading dimension ldw (nn.Linear weight)
  const float* bias;    // [N] or null
  const float* scale;   // per-row factor scale[m * scale_ld] applied after the activation, or null
  const float* colscale;  // per-column affine after the activation (BatchNorm in eval mode): v * colscale[n] + colshift[n]
  const float* colshift;
  int lda, ldw, K, act, flush, scale_ld;
  int repeat;           // > 1: the term stands for `repeat` terms r = 0..repeat-1 with W + r * w_rep, bias + r * b_rep,
  int w_rep, b_rep;     //      scale + r (the anisotropic edge filter: one [out, K] weight block per input feature)
};
struct Problem {
  Term t[MAX_TERMS];
  float* C;
  int ldc, M, N, nterms;
};
struct Batch { Problem p[MAX_PROB]; };

__device__ __forceinline__ float act_f(float v, int act) {
  switch (act) {
    case ACT_TANH: return tanhf(v);
    case ACT_RELU: return fmaxf(v, 0.f);
    case ACT_SIGMOID: return 1.0f / (1.0f + expf(-v));
    case ACT_SILU: return v / (1.0f + expf(-v));
    case ACT_ELU: return v > 0.f ? v : expm1f(v);
    default: return v;
  }
}

__device__ __forceinline__ void cp_async4(float* dst_smem, const float* src, bool valid) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
  const int n = valid ? 4 : 0;          // src-size 0: the 4 bytes are zero-filled, nothing is read
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(src), "r"(n));
}
// 16-byte copy of which only the first `bytes` (0..16) are read; the rest is zero-filled
__device__ __forceinline__ void cp_async16(float* dst_smem, const float* src, int bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(src), "r"(bytes));
}
// (lo, hi) -> one register of two bf16, round to nearest even
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// BM x BN tile, 256 threads, (BM/16) x (BN/16) outputs per thread: rows ty + 16 i, columns tx + 16 j.  K slabs of 32
// go through a GS-stage cp.async ring in the operands' own layout ([row][k], rows padded by 4 floats): 16-byte
// copy per thread and operand tile (zero-filled past K through the copy's source size; 4-byte copies when a row is not
// 16-byte aligned, e.g. the 39-wide 3D features).  The inner loop reads 4 k's of a row as one vector; with rows
// strided by 16 between a thread's outputs, the 8 lanes of a shared-memory phase hit 8 different 16-byte bank groups
// (row pitch 36 floats = 9 groups).  The small tile exists for the node-level problems (B*N rows): at 64 x 64 they are 40 CTAs.
//
// TC (64 x 64 only; models with precision = "bf16", tolerance 2e-2): the same ring, but the products run on the tensor
// cores -- mma.sync.m16n8k16, bf16 operands rounded from the fp32 tiles in registers, fp32 accumulate; 8 warps as 2 x 4,
// a warp owns 32 rows x 16 columns (2 x 2 MMAs per 16-k slab: 12 64-bit shared-memory reads, 12 packs, 4 MMAs where the
// CUDA-core loop has 32 reads and 256 FFMAs).  Biases, activations, sums over terms and outputs stay fp32.
template <int BM, int BN, int TC>
__global__ void __launch_bounds__(GT, (BM == 64 && !TC) ? 2 : (TC == 2 ? 2 : 3)) seq_gemm_kernel(const __grid_constant__ Batch bt) {
  pdl_wait();
  static_assert(!TC || (BM == 64 && BN == 64), "tensor-core variants: 64 x 64 tiles");
  // K slabs of 32 (16-wide slabs measured 2-5 % slower on the CUDA cores, 6 % in tensor-core mode: a barrier and a copy
  // group per slab either way), 3 stages for the 64 x 64 tiles, 4 for 32 x 32; dynamic shared memory (54-60 KB at 64 x 64)
  constexpr int GK = 32, GS = (BM == 64) ? 3 : 4, CPR = GK / 4;
  // row pitch: +4 floats keeps the CUDA-core loop's 128-bit reads on distinct bank groups; +8 does the same for the 64-bit
  // fragment reads of the tensor-core loop (bank = 8 g + 2 tq over a half warp)
  constexpr int TM = BM / 16, TN = BN / 16, LD = GK + (TC ? 8 : 4);
  const Problem& P = bt.p[blockIdx.z];
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  if (m0 >= P.M || n0 >= P.N) return;
  extern __shared__ __align__(16) float seq_smem[];
  float (*As)[BM][LD] = reinterpret_cast<float (*)[BM][LD]>(seq_smem);
  float (*Ws)[BN][LD] = reinterpret_cast<float (*)[BN][LD]>(seq_smem + GS * BM * LD);
  const int tid = threadIdx.x;
  // a warp covers 8 columns x 4 rows of the 16 x 16 thread grid: a 128-bit operand read touches 8 distinct W rows
  // (one 128-byte wavefront) and 4 distinct A rows (broadcast within each quarter warp) -- 8 wavefronts per 64 FFMAs
  // where a 16 x 2 warp shape needs 12
  const int tx = ((tid >> 5) & 1) * 8 + (tid & 7), ty = (tid >> 6) * 4 + ((tid >> 3) & 3);
  // TC: warp (wm, wn) of the 2 x 4 warp grid, lane = (g, tq); acc[2 mt + half][2 nt + low] is element `2 half + low` of
  // the accumulator fragment of MMA (mt, nt): row wm*32 + mt*16 + g + 8 half, column wn*16 + nt*8 + 2 tq + low
  const int lane = tid & 31, g = lane >> 2, tq = lane & 3, wm = (tid >> 5) & 1, wn = tid >> 6;
  auto row_of = [&](int i) { return TC ? wm * 32 + (i >> 1) * 16 + g + 8 * (i & 1) : ty + 16 * i; };
  auto col_of = [&](int j) { return TC ? wn * 16 + (j >> 1) * 8 + 2 * tq + (j & 1) : tx + 16 * j; };
  float sum[TM][TN], acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) { sum[i][j] = 0.f; acc[i][j] = 0.f; }

  for (int t = 0; t < P.nterms; ++t)
  for (int rep = 0; rep < (P.t[t].repeat > 1 ? P.t[t].repeat : 1); ++rep) {
    const Term& T = P.t[t];
    const int K = T.K;
    const float* Ab = T.A;
    const float* Wb = T.W + (size_t)rep * T.w_rep;
    const float* biasp = T.bias ? T.bias + (size_t)rep * T.b_rep : nullptr;
    const float* scalep = T.scale ? T.scale + rep : nullptr;
    const int lda = T.lda, ldw = T.ldw;
    const bool avec = (lda % 4 == 0) && ((((uintptr_t)Ab) & 15) == 0);
    const bool wvec = (ldw % 4 == 0) && ((((uintptr_t)Wb) & 15) == 0);
    // per-thread copy slots of a slab: chunk c = tid + 256 q covers (row c/CPR, k offset 4 (c%CPR)) of A, then of W
    constexpr int NQ = (BM + BN) * CPR / GT;
    const float* qsrc[NQ];
    unsigned qdst[NQ];
    int qlim[NQ];               // number of k's (from the slot's first one at slab 0) that may be read: 0 for rows outside
    bool qvec[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      int c = q * GT + tid;
      const bool isA = c < BM * CPR;
      if (!isA) c -= BM * CPR;
      const int row = c / CPR, kc = (c % CPR) * 4;
      const int grow = (isA ? m0 : n0) + row;
      const bool rok = grow < (isA ? P.M : P.N);
      const float* base = isA ? Ab : Wb;
      qsrc[q] = rok ? base + (size_t)grow * (isA ? lda : ldw) + kc : base;
      qdst[q] = (unsigned)__cvta_generic_to_shared(isA ? &As[0][row][kc] : &Ws[0][row][kc]);
      qlim[q] = rok ? K - kc : 0;
      qvec[q] = isA ? avec : wvec;
    }
    auto issue = [&](int kt, int stage) {
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        const int left = qlim[q] - kt * GK;                       // k's left from this chunk's first element
        const unsigned dst = qdst[q] + (unsigned)stage * (unsigned)(((q * GT + tid) < BM * CPR ? BM : BN) * LD * 4);
        if (qvec[q]) {
          const int nb = left <= 0 ? 0 : (left >= 4 ? 16 : left * 4);
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(nb > 0 ? qsrc[q] + kt * GK : qsrc[q]),
                       "r"(nb));
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int nb = left > e ? 4 : 0;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(dst + 4 * e),
                         "l"(nb > 0 ? qsrc[q] + kt * GK + e : qsrc[q]), "r"(nb));
          }
        }
      }
    };
    const int nk = (K + GK - 1) / GK;
    __syncthreads();            // the previous term's last slabs are no longer read
#pragma unroll
    for (int s = 0; s < GS - 1; ++s) {
      if (s < nk) issue(s, s);
      cp_async_commit();
    }
    for (int kt = 0; kt < nk; ++kt) {
      cp_async_wait<GS - 2>();  // slab kt has landed (for this thread's copies; the barrier covers the others)
      __syncthreads();          // ... and everybody is done with slab kt-1, whose stage is refilled next
#if !defined(SEQ_EXP_NOCOPY)      // diagnostic builds (tools/seq_exp.py): the loop without its copies / without its arithmetic
      if (kt + GS - 1 < nk) issue(kt + GS - 1, (kt + GS - 1) % GS);
#endif
      cp_async_commit();
      const int cur = kt % GS;
#if defined(SEQ_EXP_NOMATH)
      continue;
#endif
      if (TC) {
        if constexpr (TC) {
#pragma unroll
          for (int ks = 0; ks < GK / 16; ++ks) {
          // NS = 1: operands rounded to bf16.  NS = 3 (TC == 2): every fp32 operand as hi + mid + lo, three bf16 of 8
          // mantissa bits each (24 in all), and the six products whose weight is >= 2^-16 of the full one -- lo.hi,
          // hi.lo, mid.mid, mid.hi, hi.mid, hi.hi, smallest first -- accumulate in fp32: the fp32 tolerance on tensor cores
          constexpr int NS = TC == 2 ? 3 : 1;
          uint32_t afr[NS][2][4], bfr[NS][2][2];
          auto split = [&](float x, float y, uint32_t (&out)[NS]) {
            out[0] = pack_bf16x2(x, y);
            if constexpr (NS == 3) {
              const float rx = x - __uint_as_float(out[0] << 16), ry = y - __uint_as_float(out[0] & 0xffff0000u);
              out[1] = pack_bf16x2(rx, ry);
              const float sx = rx - __uint_as_float(out[1] << 16), sy = ry - __uint_as_float(out[1] & 0xffff0000u);
              out[2] = pack_bf16x2(sx, sy);
            }
          };
#pragma unroll
          for (int mt = 0; mt < 2; ++mt) {
            const float* r0 = &As[cur][wm * 32 + mt * 16 + g][16 * ks + 2 * tq];
            const float2 v[4] = {*reinterpret_cast<const float2*>(r0), *reinterpret_cast<const float2*>(r0 + 8 * LD),
                                 *reinterpret_cast<const float2*>(r0 + 8), *reinterpret_cast<const float2*>(r0 + 8 * LD + 8)};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              uint32_t o[NS];
              split(v[e].x, v[e].y, o);
#pragma unroll
              for (int q = 0; q < NS; ++q) afr[q][mt][e] = o[q];
            }
          }
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            const float* r0 = &Ws[cur][wn * 16 + nt * 8 + g][16 * ks + 2 * tq];
            const float2 v[2] = {*reinterpret_cast<const float2*>(r0), *reinterpret_cast<const float2*>(r0 + 8)};
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              uint32_t o[NS];
              split(v[e].x, v[e].y, o);
#pragma unroll
              for (int q = 0; q < NS; ++q) bfr[q][nt][e] = o[q];
            }
          }
          auto mma = [&](int mt, int nt, const uint32_t (&a)[4], const uint32_t (&b)[2]) {
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
                         "{%0, %1, %2, %3};"
                         : "+f"(acc[2 * mt][2 * nt]), "+f"(acc[2 * mt][2 * nt + 1]), "+f"(acc[2 * mt + 1][2 * nt]),
                           "+f"(acc[2 * mt + 1][2 * nt + 1])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
          };
#pragma unroll
          for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
              if constexpr (NS == 3) {
                mma(mt, nt, afr[2][mt], bfr[0][nt]);
                mma(mt, nt, afr[0][mt], bfr[2][nt]);
                mma(mt, nt, afr[1][mt], bfr[1][nt]);
                mma(mt, nt, afr[1][mt], bfr[0][nt]);
                mma(mt, nt, afr[0][mt], bfr[1][nt]);
              }
              mma(mt, nt, afr[0][mt], bfr[0][nt]);
            }
          }
        }
        continue;
      }
      float4 af[2][TM], wf[2][TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) af[0][i] = *reinterpret_cast<const float4*>(&As[cur][ty + 16 * i][0]);
#pragma unroll
      for (int j = 0; j < TN; ++j) wf[0][j] = *reinterpret_cast<const float4*>(&Ws[cur][tx + 16 * j][0]);
#pragma unroll
      for (int k4 = 0; k4 < GK / 4; ++k4) {
        const int b = k4 & 1;
        if (k4 + 1 < GK / 4) {
#pragma unroll
          for (int i = 0; i < TM; ++i) af[b ^ 1][i] = *reinterpret_cast<const float4*>(&As[cur][ty + 16 * i][4 * k4 + 4]);
#pragma unroll
          for (int j = 0; j < TN; ++j) wf[b ^ 1][j] = *reinterpret_cast<const float4*>(&Ws[cur][tx + 16 * j][4 * k4 + 4]);
        }
#pragma unroll
        for (int i = 0; i < TM; ++i)
#pragma unroll
          for (int j = 0; j < TN; ++j) {
            float v = acc[i][j];
            v = fmaf(af[b][i].x, wf[b][j].x, v);
            v = fmaf(af[b][i].y, wf[b][j].y, v);
            v = fmaf(af[b][i].z, wf[b][j].z, v);
            v = fmaf(af[b][i].w, wf[b][j].w, v);
            acc[i][j] = v;
          }
      }
    }
    cp_async_wait<0>();
#pragma unroll
    for (int i = 0; i < TM; ++i) {
      const int m = m0 + row_of(i);
      float sc = 1.0f;
      if (T.flush && scalep != nullptr && m < P.M) sc = scalep[(size_t)m * T.scale_ld];
#pragma unroll
      for (int j = 0; j < TN; ++j) {
        const int n = n0 + col_of(j);
        float v = acc[i][j];
        if (biasp != nullptr && n < P.N) v = v + biasp[n];
        if (T.flush) {
          v = act_f(v, T.act);
          if (T.colscale != nullptr && n < P.N) v = v * T.colscale[n] + T.colshift[n];
          if (scalep != nullptr) v = v * sc;
          sum[i][j] = sum[i][j] + v;
          v = 0.f;
        }
        acc[i][j] = v;
      }
    }
  }
  pdl_trigger();              // the dependents may be scheduled while the last tiles store
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + row_of(i);
    if (m >= P.M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int n = n0 + col_of(j);
      if (n < P.N) P.C[(size_t)m * P.ldc + n] = sum[i][j];
    }
  }
}

Term make_term(const float* A, int lda, const float* W, int ldw, const float* bias, int K, int act, int flush,
               const float* scale = nullptr, int scale_ld = 0) {
  Term t;
  t.A = A; t.W = W; t.bias = bias; t.scale = scale;
  t.lda = lda; t.ldw = ldw; t.K = K; t.act = act; t.flush = flush; t.scale_ld = scale_ld;
  t.colscale = nullptr; t.colshift = nullptr; t.repeat = 1; t.w_rep = 0; t.b_rep = 0;
  return t;
}

int launch_batch(const Batch& bt, int nprob, cudaStream_t st, int tc = 0) {
  int maxM = 0, maxN = 0;
  for (int i = 0; i < nprob; ++i) {
    maxM = bt.p[i].M > maxM ? bt.p[i].M : maxM;
    maxN = bt.p[i].N > maxN ? bt.p[i].N : maxN;
  }
  if (maxM == 0 || maxN == 0 || nprob == 0) return 0;
  const int64_t big = (int64_t)((maxM + 63) / 64) * ((maxN + 63) / 64) * nprob;
  // dynamic shared memory of the three instantiations (above 48 KB needs a per-device opt-in)
  constexpr int smem_tc = 3 * 128 * 40 * 4, smem_64 = 3 * 128 * 36 * 4, smem_32 = 4 * 64 * 36 * 4;
  static std::atomic<int> opted{-1};
  int dev = 0;
  cudaGetDevice(&dev);
  if (opted.load(std::memory_order_relaxed) != dev) {
    cudaError_t e = cudaFuncSetAttribute(seq_gemm_kernel<64, 64, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_tc);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(seq_gemm_kernel<64, 64, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_64);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(seq_gemm_kernel<64, 64, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_tc);
    if (e != cudaSuccess) return (int)e;
    opted.store(dev, std::memory_order_relaxed);
  }
  if (tc == 2 && big < 296 && maxM < 1024) tc = 0;   // split mode: node-level launches stay on the small CUDA-core tile
  if (tc) {           // tensor-core modes: the 64 x 64 MMA variants
    dim3 grid((maxM + 63) / 64, (maxN + 63) / 64, nprob);
    if (tc == 2) (void)seq_launch(seq_gemm_kernel<64, 64, 2>, dim3(grid), dim3(GT), smem_tc, st, bt);
    else (void)seq_launch(seq_gemm_kernel<64, 64, 1>, dim3(grid), dim3(GT), smem_tc, st, bt);
  } else if (big >= 296) {   // two resident 64x64 CTAs on every SM; below that the small tile keeps more warps in flight
    dim3 grid((maxM + 63) / 64, (maxN + 63) / 64, nprob);
    (void)seq_launch(seq_gemm_kernel<64, 64, 0>, dim3(grid), dim3(GT), smem_64, st, bt);
  } else {
    dim3 grid((maxM + 31) / 32, (maxN + 31) / 32, nprob);
    (void)seq_launch(seq_gemm_kernel<32, 32, 0>, dim3(grid), dim3(GT), smem_32, st, bt);
  }
  AE_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// elementwise pieces
// ------------------------------------------------------------------------------------------------
// m1[t][b*E+e, :] = tanh(P[b*N+recv, (2t)H + :] + P[b*N+send, (2t+1)H + :])   (":600, :609": msg_fc1 on cat[recv, send])
__global__ void seq_pair_tanh_kernel(const float* __restrict__ P, int ldp, int N, int64_t BE, int H, int t0, int nt,
                                     float* __restrict__ m1) {
  pdl_wait();
  const int H4 = H / 4;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)nt * BE * H4) return;
  const int k4 = (int)(idx % H4);
  const int64_t rest = idx / H4;
  const int64_t m = rest % BE;
  const int t = t0 + (int)(rest / BE);
  const int E = N * (N - 1);
  const int64_t b = m / E;
  const int e = (int)(m - b * E);
  const int send = e / (N - 1), q = e - send * (N - 1), recv = q + (q >= send ? 1 : 0);
  const float4 r = *reinterpret_cast<const float4*>(P + (size_t)(b * N + recv) * ldp + (size_t)(2 * t) * H + k4 * 4);
  const float4 s = *reinterpret_cast<const float4*>(P + (size_t)(b * N + send) * ldp + (size_t)(2 * t + 1) * H + k4 * 4);
  float4 o;
  o.x = tanhf(r.x + s.x); o.y = tanhf(r.y + s.y); o.z = tanhf(r.z + s.z); o.w = tanhf(r.w + s.w);
  *reinterpret_cast<float4*>(m1 + ((size_t)t * BE + m) * H + k4 * 4) = o;
}

// agg[which][b*N+j, :] = sum over senders i != j (ascending edge id) of msgs[which][b*E + e(i,j), :] / (N-1)
// (torch_scatter.scatter(reduce='mean') by receiver, ":616, :634")
// msgs1 == nullptr: one source only; add0 (optional, [B*N, H]) is added to agg0 after the division (the encoder's
// edge2node(...) + res_x, nn/seq2seq/aether.py:378).
__global__ void seq_mean_kernel(const float* __restrict__ msgs0, const float* __restrict__ msgs1, int N, int64_t B, int H,
                                float* __restrict__ agg0, float* __restrict__ agg1, const float* __restrict__ add0) {
  pdl_wait();
  const int H4 = H / 4;
  const int64_t per = B * N * H4;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (msgs1 != nullptr ? 2 : 1) * per) return;
  const int which = idx >= per;
  const int64_t r = idx - (which ? per : 0);
  const int h4 = (int)(r % H4);
  const int64_t node = r / H4;
  const int64_t b = node / N;
  const int j = (int)(node - b * N);
  const float* src = which ? msgs1 : msgs0;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  const int E = N * (N - 1);
  for (int i = 0; i < N; ++i) {
    if (i == j) continue;
    const int e = i * (N - 1) + (j < i ? j : j - 1);
    const float4 v = *reinterpret_cast<const float4*>(src + ((size_t)b * E + e) * H + h4 * 4);
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
  }
  const float cnt = (float)(N - 1);
  s.x /= cnt; s.y /= cnt; s.z /= cnt; s.w /= cnt;
  if (!which && add0 != nullptr) {
    const float4 a = *reinterpret_cast<const float4*>(add0 + (size_t)node * H + h4 * 4);
    s.x += a.x; s.y += a.y; s.z += a.z; s.w += a.w;
  }
  *reinterpret_cast<float4*>((which ? agg1 : agg0) + (size_t)node * H + h4 * 4) = s;
}

// n = tanh(inp_n + r * hidden_h(agg)); hidden' = (1 - i) n + i hidden   (":644-645")
__global__ void seq_gate_kernel(const float* __restrict__ R, const float* __restrict__ I, const float* __restrict__ NI,
                                const float* __restrict__ HH, const float* hidden_in, float* hidden_out, int64_t n) {
  pdl_wait();
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  const float r = R[idx], i = I[idx];
  const float nn = tanhf(NI[idx] + r * HH[idx]);
  hidden_out[idx] = (1.0f - i) * nn + i * hidden_in[idx];
}

// Warp per row: v[c] = A[row,:] . W[c,:] + b[c] for c < nc <= 6; with Rinv: every D-chunk rotated by Rinv[row]
// (Globalizer, nn/utils/local_to_global.py:12-13) and added to `add` (":651-652").
__global__ void seq_thin_kernel(const float* __restrict__ A, int K, const float* __restrict__ W, const float* __restrict__ b,
                                int nc, const float* __restrict__ Rinv, int D, const float* __restrict__ add,
                                float* __restrict__ out, int64_t rows) {
  pdl_wait();
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  float acc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const float4* a4 = reinterpret_cast<const float4*>(A + (size_t)row * K);     // K % 4 == 0, rows 16-byte aligned
#pragma unroll 4
  for (int k = lane; k < K / 4; k += 32) {
    const float4 a = a4[k];
#pragma unroll
    for (int c = 0; c < 6; ++c)
      if (c < nc) {
        const float4 w = reinterpret_cast<const float4*>(W + (size_t)c * K)[k];
        acc[c] = fmaf(a.x, w.x, fmaf(a.y, w.y, fmaf(a.z, w.z, fmaf(a.w, w.w, acc[c]))));
      }
  }
#pragma unroll
  for (int c = 0; c < 6; ++c)
    for (int o = 16; o > 0; o >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);
  if (lane != 0) return;
  float v[6];
  for (int c = 0; c < nc; ++c) v[c] = acc[c] + b[c];
  if (Rinv != nullptr) {
    const float* R = Rinv + (size_t)row * D * D;
    for (int c0 = 0; c0 < nc; c0 += D)
      for (int a = 0; a < D; ++a) {
        float s = R[a * D] * v[c0];
        for (int k = 1; k < D; ++k) s = s + R[a * D + k] * v[c0 + k];
        out[(size_t)row * nc + c0 + a] = add[(size_t)row * nc + c0 + a] + s;
      }
  } else {
    for (int c = 0; c < nc; ++c) out[(size_t)row * nc + c] = v[c];
  }
}

// Fourier features of the positions: [sin(p), cos(p)], p = (2 pi x) @ B   (nn/nn/fourier_feature_mapper.py:21-23)
__global__ void seq_rff_kernel(const float* __restrict__ states, int sd, int D, const float* __restrict__ Bm, int half,
                               float* __restrict__ out, int64_t rows) {
  pdl_wait();
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rows * half) return;
  const int c = (int)(idx % half);
  const int64_t row = idx / half;
  float p = (TWO_PI_F * states[row * sd]) * Bm[c];
  for (int d = 1; d < D; ++d) p = fmaf(TWO_PI_F * states[row * sd + d], Bm[(size_t)d * half + c], p);
  out[(size_t)row * 2 * half + c] = sinf(p);
  out[(size_t)row * 2 * half + half + c] = cosf(p);
}

// encoder, first Linear of mlp4 on cat[x_send, x_recv, filt] (nn/seq2seq/aether.py:379-381), node-factorised:
// out[b*E+e, :] = elu(Pn[b*N+send, 0:H] + Pn[b*N+recv, H:2H] + Pe[b*E+e, :])   (Pe carries the bias)
__global__ void seq_enc_pair_kernel(const float* __restrict__ Pn, const float* __restrict__ Pe, int N, int64_t BE, int H,
                                    float* __restrict__ out) {
  pdl_wait();
  const int H4 = H / 4;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= BE * H4) return;
  const int k4 = (int)(idx % H4);
  const int64_t m = idx / H4;
  const int E = N * (N - 1);
  const int64_t b = m / E;
  const int e = (int)(m - b * E);
  const int send = e / (N - 1), q = e - send * (N - 1), recv = q + (q >= send ? 1 : 0);
  const float4 s = *reinterpret_cast<const float4*>(Pn + (size_t)(b * N + send) * 2 * H + k4 * 4);
  const float4 r = *reinterpret_cast<const float4*>(Pn + (size_t)(b * N + recv) * 2 * H + H + k4 * 4);
  const float4 pe = *reinterpret_cast<const float4*>(Pe + (size_t)m * H + k4 * 4);
  float4 o;
  o.x = act_f(s.x + r.x + pe.x, ACT_ELU); o.y = act_f(s.y + r.y + pe.y, ACT_ELU);
  o.z = act_f(s.z + r.z + pe.z, ACT_ELU); o.w = act_f(s.w + r.w + pe.w, ACT_ELU);
  *reinterpret_cast<float4*>(out + (size_t)m * H + k4 * 4) = o;
}

// LSTM cell (nn.LSTM, gate order i, f, g, o): G[rows, 4R] holds W_ih x + b_ih + W_hh h + b_hh
__global__ void seq_lstm_cell_kernel(const float* __restrict__ G, const float* c_in, float* h_out, float* c_out, int R,
                                     int64_t rows) {
  pdl_wait();
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rows * R) return;
  const int64_t row = idx / R;
  const int j = (int)(idx - row * R);
  const float* g = G + (size_t)row * 4 * R;
  const float gi = 1.0f / (1.0f + expf(-g[j])), gf = 1.0f / (1.0f + expf(-g[R + j]));
  const float gg = tanhf(g[2 * R + j]), go = 1.0f / (1.0f + expf(-g[3 * R + j]));
  const float c = gf * c_in[idx] + gi * gg;
  c_out[idx] = c;
  h_out[idx] = go * tanhf(c);
}

// Hard gumbel-softmax sample (nn/utils/model_utils.py:57-121 with hard=True): one-hot of argmax_k(logit_k + g_k),
// g = -log(eps - log(U + eps)), first maximum on ties like torch's max; the temperature does not change an arg-max.
__global__ void seq_sample_hard_kernel(const float* __restrict__ logits, const float* __restrict__ U, float* __restrict__ out,
                                       int K, int64_t rows) {
  pdl_wait();
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  int best = 0;
  float bv = 0.f;
  for (int k = 0; k < K; ++k) {
    const float g = -logf(1e-10f - logf(U[r * K + k] + 1e-10f));
    const float v = logits[r * K + k] + g;
    if (k == 0 || v > bv) { bv = v; best = k; }
  }
  for (int k = 0; k < K; ++k) out[r * K + k] = k == best ? 1.0f : 0.0f;
}

// ------------------------------------------------------------------------------------------------
// workspace
// ------------------------------------------------------------------------------------------------
struct SeqWs {
  float *relx, *Rinv, *eattr, *P, *m1, *p1, *msgs, *pmsgs, *agg, *pagg, *gR, *gI, *gN, *gH, *o1, *o2, *f0, *f1, *f2, *field;
  float *ea, *eres, *efilt, *exin, *et1, *ex3, *ePn, *ePe, *eh1, *eemb, *eG, *ep0, *ep1;   // encoder
  size_t bytes;
};

int seq_check(const AetherSeqDims* d) {
  if (d == nullptr) return AETHER_E_BADARG;
  if (d->num_dims != 2 && d->num_dims != 3) return AETHER_E_BADARG;
  if (d->num_vars < 2 || d->num_vars > 64) return AETHER_E_BADARG;
  if (d->hidden < 4 || d->hidden % 4 != 0) return AETHER_E_UNSUPPORTED;
  if (d->n_types < 1 || d->n_types > AETHER_SEQ_MAX_TYPES) return AETHER_E_UNSUPPORTED;
  if (d->skip_first != 0 && d->skip_first != 1) return AETHER_E_BADARG;
  if (d->field_hidden < 0 || d->field_hidden % 4 != 0) return AETHER_E_UNSUPPORTED;
  if (d->rnn_hidden < 0 || d->prior_hidden < 0) return AETHER_E_BADARG;
  if (d->batch < 0) return AETHER_E_BADARG;
  const int64_t BE = d->batch * d->num_vars * (d->num_vars - 1);
  if (BE > (int64_t)1 << 30 || BE * d->hidden > ((int64_t)1 << 40)) return AETHER_E_TOO_LARGE;
  return 0;
}

SeqWs seq_layout(const AetherSeqDims* d, void* base) {
  const int D = d->num_dims, N = d->num_vars, H = d->hidden, Kt = d->n_types, Hf = d->field_hidden;
  const int O = D * (D - 1) / 2, F = 4 * D + O;
  const size_t BN = (size_t)d->batch * N, BE = BN * (N - 1);
  size_t off = 0;
  auto take = [&](size_t nfloats) {
    float* p = base ? reinterpret_cast<float*>(reinterpret_cast<char*>(base) + off) : nullptr;
    off += align_up(nfloats * sizeof(float), 256);
    return p;
  };
  SeqWs w;
  w.relx = take(BN * (3 * D + F));
  w.Rinv = take(BN * D * D);
  w.eattr = take(BE * (2 * F + 3 * D));
  w.P = take(BN * 2 * Kt * H);
  w.m1 = take((size_t)Kt * BE * H);
  w.p1 = take((size_t)Kt * BE * H);
  w.msgs = take(BE * H);
  w.pmsgs = take(BE * H);
  w.agg = take(BN * H);
  w.pagg = take(BN * H);
  w.gR = take(BN * H);
  w.gI = take(BN * H);
  w.gN = take(BN * H);
  w.gH = take(BN * H);
  w.o1 = take(BN * H);
  w.o2 = take(BN * H);
  w.f0 = take(BN * Hf);
  w.f1 = take(BN * Hf);
  w.f2 = take(BN * Hf);
  w.field = take(BN * D);
  const size_t R = (size_t)d->rnn_hidden, PH = (size_t)d->prior_hidden;
  const size_t He = R > 0 ? (size_t)Hf : 0;       // the encoder's buffers exist when an LSTM width is given
  w.ea = take(BE * He);
  w.eres = take(BN * He);
  w.efilt = take(BE * He);
  w.exin = take(BN * He);
  w.et1 = take(BN * He);
  w.ex3 = take(BN * He);
  w.ePn = take(BN * 2 * He);
  w.ePe = take(BE * He);
  w.eh1 = take(BE * He);
  w.eemb = take(BE * He);
  w.eG = take(BE * 4 * R);
  w.ep0 = take(BE * PH);
  w.ep1 = take(BE * PH);
  w.bytes = off;
  return w;
}

int seq_geometry(const AetherSeqDims* d, const float* states, const float* field, float* relx, float* Rinv, float* eattr,
                 cudaStream_t st) {
  const int N = d->num_vars;
  const int64_t total = d->batch * (int64_t)(N * (N - 1) + N);
  if (total == 0) return 0;
  const int nb = (int)ceil_div64(total, 128);
  if (d->num_dims == 2) (void)seq_launch(seq_geom_kernel<2>, dim3(nb), dim3(128), 0, st, states, field, N, d->batch, relx, Rinv, eattr);
  else (void)seq_launch(seq_geom_kernel<3>, dim3(nb), dim3(128), 0, st, states, field, N, d->batch, relx, Rinv, eattr);
  AE_LAUNCH_CHECK();
  return 0;
}

int seq_field(const AetherSeqDims* d, const AetherSeqParams* p, const float* states, float* field, const SeqWs& w,
              cudaStream_t st) {
  const int tcm = (d->flags & AETHER_SEQ_F_BF16) ? 1 : ((d->flags & AETHER_SEQ_F_SPLIT) ? 2 : 0);
  const int D = d->num_dims, Hf = d->field_hidden;
  if (Hf < 4 || p->rff_B == nullptr) return AETHER_E_BADARG;
  for (int i = 0; i < 3; ++i)
    if (p->field_w[i] == nullptr || p->field_b[i] == nullptr) return AETHER_E_BADARG;
  const int64_t BN = d->batch * d->num_vars;
  if (BN == 0) return 0;
  const int half = Hf / 2;
  (void)seq_launch(seq_rff_kernel, dim3((int)ceil_div64(BN * half, 256)), dim3(256), 0, st, states, 2 * D, D, p->rff_B, half, w.f0, BN);
  AE_LAUNCH_CHECK();
  Batch bt;
  bt.p[0].t[0] = make_term(w.f0, Hf, p->field_w[0], Hf, p->field_b[0], Hf, ACT_SILU, 1);
  bt.p[0].C = w.f1; bt.p[0].ldc = Hf; bt.p[0].M = (int)BN; bt.p[0].N = Hf; bt.p[0].nterms = 1;
  AE_CHECK(launch_batch(bt, 1, st, tcm));
  bt.p[0].t[0] = make_term(w.f1, Hf, p->field_w[1], Hf, p->field_b[1], Hf, ACT_SILU, 1);
  bt.p[0].C = w.f2;
  AE_CHECK(launch_batch(bt, 1, st, tcm));
  (void)seq_launch(seq_thin_kernel, dim3((int)ceil_div64(BN * 32, 256)), dim3(256), 0, st, w.f2, Hf, p->field_w[2], p->field_b[2], D, nullptr, D,
                                                                nullptr, field, BN);
  AE_LAUNCH_CHECK();
  return 0;
}

int seq_step(const AetherSeqDims* d, const AetherSeqParams* p, const float* states, const float* hidden_in,
             const float* edge_types, const float* field, float* states_out, float* hidden_out, const SeqWs& w,
             cudaStream_t st) {
  const int tcm = (d->flags & AETHER_SEQ_F_BF16) ? 1 : ((d->flags & AETHER_SEQ_F_SPLIT) ? 2 : 0);
  const int D = d->num_dims, N = d->num_vars, H = d->hidden, Kt = d->n_types, t0 = d->skip_first ? 1 : 0;
  const int O = D * (D - 1) / 2, F = 4 * D + O, KE = 2 * F + 3 * D, KR = 3 * D + F;
  const int64_t BN = d->batch * N, BE = BN * (N - 1);
  if (BN == 0) return 0;
  for (int t = t0; t < Kt; ++t)
    if (!p->msg_fc1_w[t] || !p->msg_fc1_b[t] || !p->msg_fc2_w[t] || !p->msg_fc2_b[t] || !p->present_msg_fc1_w[t] ||
        !p->present_msg_fc1_b[t] || !p->present_msg_fc2_w[t] || !p->present_msg_fc2_b[t])
      return AETHER_E_BADARG;
  const int nt = Kt - t0;
  const int ldp = 2 * Kt * H;

  AE_CHECK(seq_geometry(d, states, field, w.relx, w.Rinv, w.eattr, st));

  Batch bt;
  int np = 0;
  auto single = [&](const Term& t, float* C, int ldc, int M, int Nc) {
    Problem& q = bt.p[np++];
    q.t[0] = t; q.C = C; q.ldc = ldc; q.M = M; q.N = Nc; q.nterms = 1;
  };
  // batch A: node terms of msg_fc1 (bias with the receiver half) and present_msg_fc1 per type
  for (int t = t0; t < Kt; ++t) {
    if (np + 3 > MAX_PROB) {          // four active edge types: the batch goes out in two launches
      AE_CHECK(launch_batch(bt, np, st, tcm));
      np = 0;
    }
    single(make_term(hidden_in, H, p->msg_fc1_w[t], 2 * H, p->msg_fc1_b[t], H, ACT_NONE, 1), w.P + (size_t)(2 * t) * H,
           ldp, (int)BN, H);
    single(make_term(hidden_in, H, p->msg_fc1_w[t] + H, 2 * H, nullptr, H, ACT_NONE, 1), w.P + (size_t)(2 * t + 1) * H,
           ldp, (int)BN, H);
    single(make_term(w.eattr, KE, p->present_msg_fc1_w[t], KE, p->present_msg_fc1_b[t], KE, ACT_RELU, 1),
           w.p1 + (size_t)t * BE * H, H, (int)BE, H);
  }
  AE_CHECK(launch_batch(bt, np, st, tcm));
  if (nt > 0 && BE > 0) {
    const int64_t total = (int64_t)nt * BE * (H / 4);
    (void)seq_launch(seq_pair_tanh_kernel, dim3((int)ceil_div64(total, 256)), dim3(256), 0, st, w.P, ldp, N, BE, H, t0, nt, w.m1);
    AE_LAUNCH_CHECK();
  }
  // batch B: second message Linears, activation, edge-type weight, sum over types
  np = 0;
  if (nt > 0) {
    Problem& a = bt.p[np++];
    Problem& b = bt.p[np++];
    for (int t = t0; t < Kt; ++t) {
      a.t[t - t0] = make_term(w.m1 + (size_t)t * BE * H, H, p->msg_fc2_w[t], H, p->msg_fc2_b[t], H, ACT_TANH, 1,
                              edge_types + t, Kt);
      b.t[t - t0] = make_term(w.p1 + (size_t)t * BE * H, H, p->present_msg_fc2_w[t], H, p->present_msg_fc2_b[t], H,
                              ACT_RELU, 1, edge_types + t, Kt);
    }
    a.C = w.msgs; a.ldc = H; a.M = (int)BE; a.N = H; a.nterms = nt;
    b.C = w.pmsgs; b.ldc = H; b.M = (int)BE; b.N = H; b.nterms = nt;
    AE_CHECK(launch_batch(bt, np, st, tcm));
  } else {
    AE_CHECK((int)cudaMemsetAsync(w.msgs, 0, (size_t)BE * H * sizeof(float), st));
    AE_CHECK((int)cudaMemsetAsync(w.pmsgs, 0, (size_t)BE * H * sizeof(float), st));
  }
  {
    const int64_t total = 2 * BN * (H / 4);
    (void)seq_launch(seq_mean_kernel, dim3((int)ceil_div64(total, 256)), dim3(256), 0, st, w.msgs, w.pmsgs, N, d->batch, H, w.agg, w.pagg, nullptr);
    AE_LAUNCH_CHECK();
  }
  // batch C: gate pre-activations
  np = 0;
  auto gate = [&](const float* wi, const float* bi, const float* wp, const float* bp, const float* wh, int act,
                  float* C) {
    Problem& q = bt.p[np++];
    int n = 0;
    if (wi) q.t[n++] = make_term(w.relx, KR, wi, KR, bi, KR, act, 0);
    if (wp) q.t[n++] = make_term(w.pagg, H, wp, H, bp, H, act, 0);
    if (wh) q.t[n++] = make_term(w.agg, H, wh, H, nullptr, H, act, 0);
    q.t[n - 1].flush = 1;
    q.C = C; q.ldc = H; q.M = (int)BN; q.N = H; q.nterms = n;
  };
  gate(p->input_r_w, p->input_r_b, p->present_r_w, p->present_r_b, p->hidden_r_w, ACT_SIGMOID, w.gR);
  gate(p->input_i_w, p->input_i_b, p->present_i_w, p->present_i_b, p->hidden_i_w, ACT_SIGMOID, w.gI);
  gate(p->input_n_w, p->input_n_b, p->present_n_w, p->present_n_b, nullptr, ACT_NONE, w.gN);
  gate(nullptr, nullptr, nullptr, nullptr, p->hidden_h_w, ACT_NONE, w.gH);
  AE_CHECK(launch_batch(bt, np, st, tcm));
  (void)seq_launch(seq_gate_kernel, dim3((int)ceil_div64(BN * H, 256)), dim3(256), 0, st, w.gR, w.gI, w.gN, w.gH, hidden_in, hidden_out, BN * H);
  AE_LAUNCH_CHECK();
  // out MLP
  np = 0;
  single(make_term(hidden_out, H, p->out_w[0], H, p->out_b[0], H, ACT_RELU, 1), w.o1, H, (int)BN, H);
  AE_CHECK(launch_batch(bt, np, st, tcm));
  np = 0;
  single(make_term(w.o1, H, p->out_w[1], H, p->out_b[1], H, ACT_RELU, 1), w.o2, H, (int)BN, H);
  AE_CHECK(launch_batch(bt, np, st, tcm));
  (void)seq_launch(seq_thin_kernel, dim3((int)ceil_div64(BN * 32, 256)), dim3(256), 0, st, w.o2, H, p->out_w[2], p->out_b[2], 2 * D, w.Rinv, D,
                                                                states, states_out, BN);
  AE_LAUNCH_CHECK();
  return 0;
}

int seq_params_ok(const AetherSeqParams* p) {
  if (p == nullptr) return AETHER_E_BADARG;
  if (!p->hidden_r_w || !p->hidden_i_w || !p->hidden_h_w || !p->present_r_w || !p->present_r_b || !p->present_i_w ||
      !p->present_i_b || !p->present_n_w || !p->present_n_b || !p->input_r_w || !p->input_r_b || !p->input_i_w ||
      !p->input_i_b || !p->input_n_w || !p->input_n_b)
    return AETHER_E_BADARG;
  for (int i = 0; i < 3; ++i)
    if (!p->out_w[i] || !p->out_b[i]) return AETHER_E_BADARG;
  return 0;
}

int seq_encoder(const AetherSeqDims* d, const AetherSeqEncParams* p, const float* states, const float* field,
                const float* h_in, const float* c_in, float* logits, float* h_out, float* c_out, const SeqWs& w,
                cudaStream_t st) {
  const int tcm = (d->flags & AETHER_SEQ_F_BF16) ? 1 : ((d->flags & AETHER_SEQ_F_SPLIT) ? 2 : 0);
  const int D = d->num_dims, N = d->num_vars, H = d->field_hidden, R = d->rnn_hidden, PH = d->prior_hidden, Kt = d->n_types;
  const int O = D * (D - 1) / 2, F = 4 * D + O, KE = 2 * F + 3 * D, KR = 3 * D + F, KP = D + O;
  const int64_t BN = d->batch * N, BE = BN * (N - 1);
  if (H < 4 || R < 1 || p->n_prior_layers < 1 || p->n_prior_layers > 4 || (p->n_prior_layers > 1 && PH < 1))
    return AETHER_E_BADARG;
  if (p->pos_offset < 0 || p->pos_offset + KP > F) return AETHER_E_BADARG;
  if (!p->filter_w0 || !p->filter_b0 || !p->filter_w2 || !p->filter_b2 || !p->res1_w || !p->res1_b || !p->lstm_w_ih ||
      !p->lstm_w_hh || !p->lstm_b_ih || !p->lstm_b_hh)
    return AETHER_E_BADARG;
  for (int i = 0; i < 2; ++i)
    if (!p->mlp3_w[i] || !p->mlp3_b[i] || !p->mlp4_w[i] || !p->mlp4_b[i]) return AETHER_E_BADARG;
  for (int i = 0; i < p->n_prior_layers; ++i)
    if (!p->prior_w[i] || !p->prior_b[i]) return AETHER_E_BADARG;
  if (BN == 0) return 0;

  AE_CHECK(seq_geometry(d, states, field, w.relx, w.Rinv, w.eattr, st));
  Batch bt;
  int np = 0;
  auto add = [&](const Term& t, float* C, int ldc, int64_t M, int Nc) -> Problem& {
    Problem& q = bt.p[np++];
    q.t[0] = t; q.C = C; q.ldc = ldc; q.M = (int)M; q.N = Nc; q.nterms = 1;
    return q;
  };
  // filter generator's first Linear on the edge positions (a column range of edge_attr) and the residual Linear
  add(make_term(w.eattr + p->pos_offset, KE, p->filter_w0, KP, p->filter_b0, KP, ACT_ELU, 1), w.ea, H, BE, H);
  add(make_term(w.relx, KR, p->res1_w, KR, p->res1_b, KR, ACT_NONE, 1), w.eres, H, BN, H);
  AE_CHECK(launch_batch(bt, np, st, tcm));
  // anisotropic filter (nn/nn/anisotropic_filter.py:35-41): filt[e,:] = sum_f attr[e,f] (a[e,:] W2_f^T + b2_f), W2_f the
  // f-th block of H rows of edge_filter.2 -- one repeated term, the [E, Fin*H] weight tensor is never formed
  np = 0;
  {
    Term t = make_term(w.ea, H, p->filter_w2, H, p->filter_b2, H, ACT_NONE, 1, w.eattr, KE);
    t.repeat = KE; t.w_rep = H * H; t.b_rep = H;
    add(t, w.efilt, H, BE, H);
  }
  AE_CHECK(launch_batch(bt, np, st, tcm));
  (void)seq_launch(seq_mean_kernel, dim3((int)ceil_div64(BN * (H / 4), 256)), dim3(256), 0, st, w.efilt, nullptr, N, d->batch, H, w.exin, nullptr,
                                                                     w.eres);
  AE_LAUNCH_CHECK();
  np = 0;
  add(make_term(w.exin, H, p->mlp3_w[0], H, p->mlp3_b[0], H, ACT_ELU, 1), w.et1, H, BN, H);
  AE_CHECK(launch_batch(bt, np, st, tcm));
  np = 0;
  {
    Problem& q = add(make_term(w.et1, H, p->mlp3_w[1], H, p->mlp3_b[1], H, ACT_ELU, 1), w.ex3, H, BN, H);
    q.t[0].colscale = p->mlp3_bn_scale; q.t[0].colshift = p->mlp3_bn_shift;
  }
  AE_CHECK(launch_batch(bt, np, st, tcm));
  np = 0;   // mlp4's first Linear [H, 3H] on cat[x_send, x_recv, filt]: the two node parts per particle, the edge part per pair
  add(make_term(w.ex3, H, p->mlp4_w[0], 3 * H, nullptr, H, ACT_NONE, 1), w.ePn, 2 * H, BN, H);
  add(make_term(w.ex3, H, p->mlp4_w[0] + H, 3 * H, nullptr, H, ACT_NONE, 1), w.ePn + H, 2 * H, BN, H);
  add(make_term(w.efilt, H, p->mlp4_w[0] + 2 * H, 3 * H, p->mlp4_b[0], H, ACT_NONE, 1), w.ePe, H, BE, H);
  AE_CHECK(launch_batch(bt, np, st, tcm));
  if (BE > 0) {
    (void)seq_launch(seq_enc_pair_kernel, dim3((int)ceil_div64(BE * (H / 4), 256)), dim3(256), 0, st, w.ePn, w.ePe, N, BE, H, w.eh1);
    AE_LAUNCH_CHECK();
  }
  np = 0;
  {
    Problem& q = add(make_term(w.eh1, H, p->mlp4_w[1], H, p->mlp4_b[1], H, ACT_ELU, 1), w.eemb, H, BE, H);
    q.t[0].colscale = p->mlp4_bn_scale; q.t[0].colshift = p->mlp4_bn_shift;
  }
  AE_CHECK(launch_batch(bt, np, st, tcm));
  np = 0;   // LSTM gates
  {
    Problem& q = bt.p[np++];
    q.t[0] = make_term(w.eemb, H, p->lstm_w_ih, H, p->lstm_b_ih, H, ACT_NONE, 0);
    q.t[1] = make_term(h_in, R, p->lstm_w_hh, R, p->lstm_b_hh, R, ACT_NONE, 1);
    q.C = w.eG; q.ldc = 4 * R; q.M = (int)BE; q.N = 4 * R; q.nterms = 2;
  }
  AE_CHECK(launch_batch(bt, np, st, tcm));
  if (BE > 0) {
    (void)seq_launch(seq_lstm_cell_kernel, dim3((int)ceil_div64(BE * R, 256)), dim3(256), 0, st, w.eG, c_in, h_out, c_out, R, BE);
    AE_LAUNCH_CHECK();
  }
  const float* cur = h_out;
  int kin = R;
  for (int i = 0; i < p->n_prior_layers; ++i) {
    const bool last = i == p->n_prior_layers - 1;
    float* dst = last ? logits : (i & 1 ? w.ep1 : w.ep0);
    const int nout = last ? Kt : PH;
    np = 0;
    add(make_term(cur, kin, p->prior_w[i], kin, p->prior_b[i], kin, last ? ACT_NONE : ACT_ELU, 1), dst, nout, BE, nout);
    AE_CHECK(launch_batch(bt, np, st, tcm));
    cur = dst;
    kin = nout;
  }
  return 0;
}

}  // namespace
}  // namespace aether

using namespace aether;

extern "C" size_t aether_seq_workspace_bytes(const AetherSeqDims* dims) {
  if (seq_check(dims) != 0) return 0;
  return seq_layout(dims, nullptr).bytes + 256;
}

extern "C" int aether_seq_localizer(const AetherSeqDims* dims, const float* states, const float* field, float* rel_feat,
                                    float* Rinv, float* edge_attr, void* stream) {
  AE_CHECK(seq_check(dims));
  if (!states || !field || !rel_feat || !Rinv || !edge_attr) return AETHER_E_BADARG;
  return seq_geometry(dims, states, field, rel_feat, Rinv, edge_attr, (cudaStream_t)stream);
}

static int seq_ws(const AetherSeqDims* dims, void* workspace, size_t workspace_bytes, SeqWs* w) {
  AE_CHECK(seq_check(dims));
  if (workspace == nullptr) return AETHER_E_WORKSPACE;
  void* base = reinterpret_cast<void*>(align_up((size_t)(uintptr_t)workspace, 256));
  *w = seq_layout(dims, base);
  if ((size_t)((char*)base - (char*)workspace) + w->bytes > workspace_bytes) return AETHER_E_WORKSPACE;
  return 0;
}

extern "C" int aether_seq_predict_field(const AetherSeqDims* dims, const AetherSeqParams* params, const float* states,
                                        float* field, void* workspace, size_t workspace_bytes, void* stream) {
  SeqWs w;
  AE_CHECK(seq_ws(dims, workspace, workspace_bytes, &w));
  if (!params || !states || !field) return AETHER_E_BADARG;
  return seq_field(dims, params, states, field, w, (cudaStream_t)stream);
}

extern "C" int aether_seq_decoder_step(const AetherSeqDims* dims, const AetherSeqParams* params, const float* states,
                                       const float* hidden_in, const float* edge_types, const float* field,
                                       float* states_out, float* hidden_out, void* workspace, size_t workspace_bytes,
                                       void* stream) {
  SeqWs w;
  AE_CHECK(seq_ws(dims, workspace, workspace_bytes, &w));
  AE_CHECK(seq_params_ok(params));
  if (!states || !hidden_in || !edge_types || !field || !states_out || !hidden_out) return AETHER_E_BADARG;
  return seq_step(dims, params, states, hidden_in, edge_types, field, states_out, hidden_out, w, (cudaStream_t)stream);
}

extern "C" int aether_seq_rollout_step(const AetherSeqDims* dims, const AetherSeqParams* params, const float* states,
                                       const float* hidden_in, const float* edge_types, float* states_out,
                                       float* hidden_out, float* field_out, void* workspace, size_t workspace_bytes,
                                       void* stream) {
  SeqWs w;
  AE_CHECK(seq_ws(dims, workspace, workspace_bytes, &w));
  AE_CHECK(seq_params_ok(params));
  if (!states || !hidden_in || !edge_types || !states_out || !hidden_out) return AETHER_E_BADARG;
  float* field = field_out ? field_out : w.field;
  AE_CHECK(seq_field(dims, params, states, field, w, (cudaStream_t)stream));
  return seq_step(dims, params, states, hidden_in, edge_types, field, states_out, hidden_out, w, (cudaStream_t)stream);
}

extern "C" int aether_seq_encoder_step(const AetherSeqDims* dims, const AetherSeqEncParams* params, const float* states,
                                       const float* field, const float* h_in, const float* c_in, float* logits,
                                       float* h_out, float* c_out, void* workspace, size_t workspace_bytes, void* stream) {
  SeqWs w;
  AE_CHECK(seq_ws(dims, workspace, workspace_bytes, &w));
  if (!params || !states || !field || !h_in || !c_in || !logits || !h_out || !c_out) return AETHER_E_BADARG;
  if (dims->rnn_hidden < 1) return AETHER_E_BADARG;
  return seq_encoder(dims, params, states, field, h_in, c_in, logits, h_out, c_out, w, (cudaStream_t)stream);
}

extern "C" int aether_seq_sample_hard(const float* logits, const float* uniform, float* onehot, int64_t rows, int32_t n_types,
                                      void* stream) {
  if (!logits || !uniform || !onehot || rows < 0 || n_types < 1) return AETHER_E_BADARG;
  if (rows == 0) return 0;
  (void)seq_launch(seq_sample_hard_kernel, dim3((unsigned)ceil_div64(rows, 256)), dim3(256), 0, (cudaStream_t)stream, logits,
                   uniform, onehot, (int)n_types, rows);
  AE_LAUNCH_CHECK();
  return 0;
}
