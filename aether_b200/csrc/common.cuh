// Shared declarations of the aether_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <atomic>
#include <stdint.h>

#include "../../include/aether_b200.h"

#define AE_CHECK(expr)                              \
  do {                                              \
    int _e = (int)(expr);                           \
    if (_e != 0) return _e;                         \
  } while (0)

#define AE_LAUNCH_CHECK()                           \
  do {                                              \
    ::aether::g_launch_count.fetch_add(1, std::memory_order_relaxed); \
    cudaError_t _e = cudaPeekAtLastError();         \
    if (_e != cudaSuccess) return (int)_e;          \
  } while (0)

namespace aether {

// prof.cu: number of kernels launched by this library (every launch site runs AE_LAUNCH_CHECK)
extern std::atomic<unsigned long long> g_launch_count;
// Optional CUDA-event timing of a group of launches; a no-op unless aether_profile_enable(1).
enum ProfCat { PROF_EDGE_MLP_FWD = 0, PROF_EDGE_BWD_DATA = 1, PROF_EDGE_WGRAD = 2, PROF_SEGMENT = 3,
               PROF_EDGE_FEAT = 4, PROF_NODE = 5, PROF_SEGMEAN_FWD = 6, PROF_NUM = 7 };
struct ProfScope {
  ProfScope(int cat, cudaStream_t st);
  ~ProfScope();
  cudaStream_t st_; long idx_;
};

static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Edge-feature geometry.  Fe = 7D + O + 2 (nn/state2state/locs/locs.py:170-178).
__host__ __device__ inline int num_euler(int D) { return D * (D - 1) / 2; }
__host__ __device__ inline int num_edge_feat(int D) { return 7 * D + num_euler(D) + 2; }
// leading dimension of the stored feature matrix (rows 16-byte aligned)
__host__ __device__ inline int edge_feat_ld(int D) { return (num_edge_feat(D) + 3) / 4 * 4; }

__device__ __forceinline__ float silu_f(float z) { return z / (1.0f + expf(-z)); }
// d silu / dz, the form ATen's silu_backward uses: s * (1 + z * (1 - s))
__device__ __forceinline__ float dsilu_f(float z) {
  float s = 1.0f / (1.0f + expf(-z));
  return s * (1.0f + z * (1.0f - s));
}

// Planar distance of the kNN builders (graph.cu, small.cu): one definition with the roundings spelled out, so that both
// kernels pick bit-identical neighbours whatever the file's contraction flags are.
__device__ __forceinline__ float knn_dist(float xi, float yi, float xj, float yj) {
  const float dx = __fsub_rn(xi, xj), dy = __fsub_rn(yi, yj);
  return sqrtf(__fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

// ---- generic fp32 GEMM family (gemm_f32.cu) ---------------------------------
enum GemmLayout { GEMM_NT = 0, GEMM_NN = 1, GEMM_TN = 2 };
enum LoadTf { TF_NONE = 0, TF_SILU = 1, TF_SILU_MASK = 2 };

struct GemmArgs {
  // C[M,N] = epi( sum_k Aop(m,k) * Bop(k,n) )
  //   NT: Aop(m,k)=tf(A[m*lda+k])   Bop(k,n)=B[n*ldb+k]      (y = x W^T)
  //   NN: Aop(m,k)=A[m*lda+k]       Bop(k,n)=B[k*ldb+n]      (dx = dy W)
  //   TN: Aop(m,k)=A[k*lda+m]       Bop(k,n)=tf(B[k*ldb+n])  (dW = dy^T x), K = rows reduced over
  const float* A = nullptr; int64_t lda = 0;
  const float* B = nullptr; int64_t ldb = 0;
  float* C = nullptr; int64_t ldc = 0;
  int64_t M = 0; int N = 0; int64_t K = 0;
  int tf = TF_NONE;                       // transform on the activations operand (A for NT, B for TN)
  const float* tf_mask = nullptr;         // [rows, cols] same ld as the transformed operand
  // epilogue (NT/NN only; applied in this order)
  const float* bias = nullptr;            // + bias[n]
  const float* gat = nullptr; int64_t ldg = 0;  // + gat[idx_a[m]*ldg + n] + gat[idx_b[m]*ldg + goff_b + n]
  const int32_t* idx_a = nullptr; const int32_t* idx_b = nullptr; int64_t goff_b = 0;
  const float* res = nullptr; int64_t ldr = 0;  // + res[m*ldr + n]
  const float* dsilu = nullptr; int64_t ldd = 0;  // * silu'(dsilu[m*ldd+n])
  const float* emask = nullptr; int64_t ldm = 0;  // * emask[m*ldm+n]
  int accumulate = 0;                     // C += result
  // TN only: deterministic split over K
  float* partial = nullptr; size_t partial_floats = 0;
};

int gemm_f32(int layout, const GemmArgs& a, cudaStream_t st);
// thin.cu: the same contract for shapes with one thin side; AETHER_E_UNSUPPORTED for every other shape
int gemm_f32_thin(int layout, const GemmArgs& a, cudaStream_t st);
size_t gemm_tn_partial_floats(int64_t M, int N, int64_t K);

// column sums: out[n] = sum_m tf? A[m,n]  (deterministic two-stage)
int colsum_f32(const float* A, int64_t lda, int64_t M, int N, float* out, float* partial, size_t partial_floats,
               cudaStream_t st);
size_t colsum_partial_floats(int64_t M, int N);

// ---- geometry / node / edge kernels (geom.cu, compiled with -fmad=false) ----
struct NodeBufs {
  const float* x; const float* vel; const float* charges;
  float* fin;      // [Nn, 2D+16] field-network input
  float* f;        // [Nn, D]   field
  float* R;        // [Nn, D*D]
  float* x0;       // [Nn, 3D]  rel_feat
};
int launch_field_input(int D, int64_t Nn, const float* x, const float* vel, const float* charges,
                       const float* emb /*[3,16]*/, float* fin, cudaStream_t st);
// the whole field MLP in one launch (field.cu): fin, the two pre-activations [Nn,32] and the field [Nn,D]
int launch_field_mlp_fwd(int D, int64_t Nn, const float* x, const float* vel, const float* charges, const float* emb,
                         const float* W0, const float* b0, const float* W2, const float* b2, const float* W4, const float* b4,
                         float* fin, float* z1, float* z2, float* f, cudaStream_t st);
size_t field_mlp_bwd_partial_floats(int D, int64_t Nn);
int launch_field_mlp_bwd(int D, int64_t Nn, const float* gf, const float* fin, const float* z1, const float* z2, const float* W0,
                         const float* W2, const float* W4, float* gfin, float* dW0, float* db0, float* dW2, float* db2,
                         float* dW4, float* db4, float* partial, size_t partial_floats, cudaStream_t st);
int launch_node_frame(int D, int64_t Nn, const float* vel, const float* f, float* R, float* x0, cudaStream_t st);
int launch_edge_features(int D, int64_t E, const int32_t* send, const int32_t* recv, const int32_t* perm,
                         const float* x, const float* vel, const float* f, const float* R, const float* x0,
                         const float* edge_attr_orig, float* feat, int ldf, cudaStream_t st);
// bf16 path: features as [E, 64] bf16 = [hi(32) | lo(32)] (A operand of the layer-1 tcgen05 GEMM)
int launch_edge_features_hilo(int D, int64_t E, const int32_t* send, const int32_t* recv, const int32_t* perm,
                              const float* x, const float* vel, const float* f, const float* R, const float* x0,
                              const float* edge_attr_orig, void* feat16, cudaStream_t st);
int launch_globalize(int D, int64_t Nn, const float* x, const float* R, const float* pred, float* out, cudaStream_t st);
int launch_caller_prep(int D, const float* x, const float* vel, const float* q, const int64_t* send,
                       const int64_t* recv, int64_t E, int64_t Nn, float* h, float* ea, cudaStream_t st);
// backward of the geometry (geom_bwd.cu)
int launch_globalize_bwd(int D, int64_t Nn, const float* R, const float* pred, const float* gout, float* gpred,
                         float* gR /*[Nn,D*D] overwritten*/, cudaStream_t st);
int launch_edge_features_bwd(int D, int64_t E, const int32_t* send, const int32_t* recv, const float* x,
                             const float* vel, const float* f, const float* R, const float* gfeat, int ldf,
                             float* gsend /*[E, GS]*/, float* grecv /*[E, GR]*/, cudaStream_t st);
int launch_node_frame_bwd(int D, int64_t Nn, const float* vel, const float* f, const float* R, const float* gx0,
                          const float* gsend_sum, const float* grecv_sum, const float* gR_glob, const float* gout,
                          float* gf, float* gx, float* gvel, cudaStream_t st);
// gx[n,:] += gfin[n,0:D] ; gvel[n,:] += gfin[n,D:2D]   (either output may be null)
int launch_add_field_input_grads(int D, int64_t Nn, const float* gfin, float* gx, float* gvel, cudaStream_t st);
__host__ __device__ inline int geom_gs(int D) { return 3 * D + D * D; }   // per-edge sender-side grads
__host__ __device__ inline int geom_gr(int D) { return 3 * D + D * D; }   // per-edge receiver-side grads

// ---- segment kernels (segment.cu) --------------------------------------------
// y[n,:] = base[n,:] + mean_{k in seg(n)} silu(src[k,:])   (base may be null)
int launch_segmean_silu(int64_t Nn, int H, const int32_t* rowptr, const float* src, int64_t lds, const float* base,
                        int64_t ldb, float* y, int64_t ldy, cudaStream_t st);
// out[n,:] = sum_{k in [ptr[n],ptr[n+1])} src[(list? list[k] : k), :]
int launch_segsum(int64_t Nn, int W, const int32_t* ptr, const int32_t* list, const float* src, int64_t lds,
                  float* out, int64_t ldo, int accumulate, cudaStream_t st);
// g[k,:] = (ge? ge[k,:] : 0 + gy[recv[k],:]/max(deg,1)) * silu'(pre2[k,:])
int launch_edge_dpre2(int64_t E, int H, const int32_t* recv, const int32_t* rowptr, const float* ge, const float* gy,
                      const float* pre2, float* g, cudaStream_t st);
// embedding gradient: gemb[r, :] = sum_{n: idx(n)==r} gfin[n, 2D:2D+16]
int launch_embedding_bwd(int D, int64_t Nn, const float* charges, const float* gfin, float* gemb, float* partial,
                         size_t partial_floats, cudaStream_t st);
int launch_fill(float* p, size_t n, float v, cudaStream_t st);

// ---- bf16 edge kernels (edge_bf16.cu): [E, W] tensors stored as bf16, W in {64, 128, 256} ----------
int launch_segmean_silu_bf16(int64_t Nn, int H, const int32_t* rowptr, const void* src, int64_t lds, const float* base,
                             int64_t ldb, float* y, int64_t ldy, cudaStream_t st, void* y16 = nullptr, int64_t ldy16 = 0);
int launch_segsum_bf16(int64_t Nn, int W, const int32_t* ptr, const int32_t* list, const void* src, int64_t lds,
                       float* out, int64_t ldo, cudaStream_t st, void* out16 = nullptr, int64_t ldo16 = 0);
int launch_edge_dpre2_bf16(int64_t E, int H, const int32_t* recv, const int32_t* rowptr, const void* ge, const float* gy,
                           const void* pre2, void* g, cudaStream_t st);
// out16[n,:] = bf16(src[n,:] / max(deg(n), 1))   (the mean's backward, gathered per edge by the next GEMM)
int launch_node_mean_scale_bf16(int64_t Nn, int H, const int32_t* rowptr, const float* src, void* out16, cudaStream_t st);
int colsum_bf16(const void* A, int64_t lda, int64_t M, int N, float* out, float* partial, size_t partial_floats,
                cudaStream_t st);
size_t colsum_bf16_partial_floats(int64_t M, int N);

// ---- graph kernels (graph.cu) --------------------------------------------------
}  // namespace aether
