// Integer path: receiver-sorted CSR of an edge list (stable, bit-exact with
// argsort(recv, stable) / bincount / cumsum), the sender-grouped view used by backward, and the
// reference's edge-list builders on the device
// (experiments/lorentz/dataset4newton.py:54-61,84-94; experiments/ind/single_ind_data.py:186-217).
#include "common.cuh"

namespace aether {

namespace {

template <typename T>
__global__ void count_kernel(const T* __restrict__ key, int64_t E, int64_t Nn, int32_t* cnt, int32_t* status) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const long long k = (long long)key[e];
  if (k < 0 || k >= Nn) { atomicOr(status, 1); return; }
  atomicAdd(&cnt[k], 1);    // integer histogram: the result does not depend on arrival order
}

// Exclusive scan of cnt[0..Nn) into ptr[0..Nn]; single block, chunked with a running carry.
__global__ void scan_kernel(const int32_t* __restrict__ cnt, int64_t Nn, int32_t* __restrict__ ptr) {
  __shared__ int32_t warp_tot[32];
  __shared__ int32_t carry_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < Nn; base += blockDim.x) {
    const int64_t i = base + tid;
    const int32_t v = i < Nn ? cnt[i] : 0;
    int32_t s = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t t = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += t;
    }
    if (lane == 31) warp_tot[wid] = s;
    __syncthreads();
    if (wid == 0) {
      int32_t w = lane < (int)(blockDim.x >> 5) ? warp_tot[lane] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int32_t t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      warp_tot[lane] = w;   // inclusive over warps
    }
    __syncthreads();
    const int32_t carry = carry_s;
    const int32_t excl = carry + (wid ? warp_tot[wid - 1] : 0) + s - v;
    if (i < Nn) ptr[i] = excl;
    __syncthreads();
    if (tid == blockDim.x - 1) carry_s = carry + warp_tot[(blockDim.x >> 5) - 1];
    __syncthreads();
  }
  if (tid == 0) ptr[Nn] = carry_s;
}

template <typename T>
__global__ void place_kernel(const T* __restrict__ key, int64_t E, int64_t Nn, const int32_t* __restrict__ ptr,
                             int32_t* fill, int32_t* __restrict__ tmp) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const long long k = (long long)key[e];
  if (k < 0 || k >= Nn) return;
  const int32_t pos = ptr[k] + atomicAdd(&fill[k], 1);   // unordered inside the segment; fixed next
  tmp[pos] = (int32_t)e;
}

// Sort each segment ascending by edge id (= stable order).  Rank by counting: ids are distinct, so
// rank(i) = #{j : id_j < id_i}.  One warp per segment for degree <= BIG, one block otherwise.
constexpr int BIG_DEG = 2048;

__global__ void segsort_warp_kernel(int64_t Nn, const int32_t* __restrict__ ptr, const int32_t* __restrict__ tmp,
                                    int32_t* __restrict__ out) {
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (n >= Nn) return;
  const int b = ptr[n], e = ptr[n + 1], deg = e - b;
  if (deg > BIG_DEG) return;
  if (deg <= 32) {
    const int32_t mine = lane < deg ? tmp[b + lane] : 0x7fffffff;
    int rank = 0;
    for (int j = 0; j < deg; ++j) rank += __shfl_sync(0xffffffffu, mine, j) < mine;
    if (lane < deg) out[b + rank] = mine;
    return;
  }
  for (int i = lane; i < deg; i += 32) {
    const int32_t mine = tmp[b + i];
    int rank = 0;
    for (int j = 0; j < deg; ++j) rank += tmp[b + j] < mine;
    out[b + rank] = mine;
  }
}

__global__ void segsort_block_kernel(int64_t Nn, const int32_t* __restrict__ ptr, const int32_t* __restrict__ tmp,
                                     int32_t* __restrict__ out) {
  const int64_t n = blockIdx.x;
  const int b = ptr[n], e = ptr[n + 1], deg = e - b;
  if (deg <= BIG_DEG) return;
  __shared__ int32_t tile[1024];
  for (int i0 = 0; i0 < deg; i0 += blockDim.x) {
    const int i = i0 + threadIdx.x;
    const int32_t mine = i < deg ? tmp[b + i] : 0x7fffffff;
    int rank = 0;
    for (int j0 = 0; j0 < deg; j0 += 1024) {
      __syncthreads();
      for (int t = threadIdx.x; t < 1024; t += blockDim.x) tile[t] = j0 + t < deg ? tmp[b + j0 + t] : 0x7fffffff;
      __syncthreads();
      const int lim = deg - j0 < 1024 ? deg - j0 : 1024;
      for (int j = 0; j < lim; ++j) rank += tile[j] < mine;
    }
    if (i < deg) out[b + rank] = mine;
  }
}

template <typename T>
__global__ void gather_kernel(int64_t E, const int32_t* __restrict__ perm, const T* __restrict__ a,
                              const T* __restrict__ b, int32_t* __restrict__ oa, int32_t* __restrict__ ob) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const int32_t p = perm[e];
  oa[e] = (int32_t)a[p];
  ob[e] = (int32_t)b[p];
}

__global__ void fc_edges_kernel(int64_t n, int64_t B, int64_t* __restrict__ send, int64_t* __restrict__ recv) {
  const int64_t per = n * (n - 1);
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= per * B) return;
  const int64_t g = t / per, r = t % per;
  const int64_t i = r / (n - 1);
  int64_t j = r % (n - 1);
  if (j >= i) ++j;                       // sender-major, j != i in ascending order
  send[t] = i + g * n;
  recv[t] = j + g * n;
}

// One block per scene (n <= 1024 agents), one thread per query agent.  k smallest distances in
// ascending order; ties resolved toward the lower index.  Distances as cdist computes them
// (sqrt of the summed squared differences) on the first two coordinates.
__global__ void knn_kernel(const float* __restrict__ pos, int stride, const int32_t* __restrict__ node_off,
                           const int64_t* __restrict__ edge_off, int k, int64_t* __restrict__ send,
                           int64_t* __restrict__ recv) {
  extern __shared__ float sp[];
  const int s = blockIdx.x;
  const int n0 = node_off[s], n = node_off[s + 1] - n0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    sp[2 * i] = pos[(size_t)(n0 + i) * stride];
    sp[2 * i + 1] = pos[(size_t)(n0 + i) * stride + 1];
  }
  __syncthreads();
  if (n <= 1) return;
  const int kk = k < n - 1 ? k : n - 1;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float xi = sp[2 * i], yi = sp[2 * i + 1];
    float last_d = -1.0f;
    int last_j = -1;
    const int64_t base = edge_off[s] + (int64_t)i * kk;
    for (int t = 0; t < kk; ++t) {
      float best = INFINITY;
      int bj = -1;
      for (int j = 0; j < n; ++j) {
        if (j == i) continue;
        const float dx = xi - sp[2 * j], dy = yi - sp[2 * j + 1];
        const float d = sqrtf(dx * dx + dy * dy);
        const bool after = d > last_d || (d == last_d && j > last_j);
        if (after && (d < best || (d == best && j < bj) || bj < 0)) { best = d; bj = j; }
      }
      send[base + t] = n0 + i;
      recv[base + t] = n0 + bj;
      last_d = best;
      last_j = bj;
    }
  }
}

inline unsigned nblk(int64_t n, int b) { return (unsigned)((n + b - 1) / b); }

struct Scratch { int32_t* cnt; int32_t* fill; int32_t* tmp; };

size_t scratch_bytes(int64_t Nn, int64_t E) {
  return align_up((size_t)(Nn + 1) * 4, 256) * 2 + align_up((size_t)(E > 0 ? E : 1) * 4, 256);
}

template <typename T>
int csr_by_key(const T* key, int64_t E, int64_t Nn, int32_t* ptr, int32_t* order, int32_t* status, Scratch s,
               cudaStream_t st) {
  AE_CHECK(cudaMemsetAsync(s.cnt, 0, (size_t)(Nn + 1) * 4, st));
  AE_CHECK(cudaMemsetAsync(s.fill, 0, (size_t)(Nn + 1) * 4, st));
  if (E > 0) {
    count_kernel<T><<<nblk(E, 256), 256, 0, st>>>(key, E, Nn, s.cnt, status);
    AE_LAUNCH_CHECK();
  }
  scan_kernel<<<1, 1024, 0, st>>>(s.cnt, Nn, ptr);
  AE_LAUNCH_CHECK();
  if (E > 0) {
    place_kernel<T><<<nblk(E, 256), 256, 0, st>>>(key, E, Nn, ptr, s.fill, s.tmp);
    AE_LAUNCH_CHECK();
    segsort_warp_kernel<<<nblk(Nn * 32, 256), 256, 0, st>>>(Nn, ptr, s.tmp, order);
    AE_LAUNCH_CHECK();
    segsort_block_kernel<<<(unsigned)Nn, 256, 0, st>>>(Nn, ptr, s.tmp, order);
    AE_LAUNCH_CHECK();
  }
  return 0;
}

}  // namespace
}  // namespace aether

using namespace aether;

extern "C" size_t aether_graph_scratch_bytes(int64_t n_nodes, int64_t n_edges) {
  return scratch_bytes(n_nodes, n_edges);
}

extern "C" int aether_graph_prepare(const int64_t* send, const int64_t* recv, int64_t E, int64_t Nn,
                                    AetherGraph* g, int32_t* status, void* scratch, size_t bytes, void* stream) {
  if (!g || !status || E < 0 || Nn < 0 || E >= (int64_t)1 << 31 || Nn >= (int64_t)1 << 31) return AETHER_E_BADARG;
  if (bytes < scratch_bytes(Nn, E) || !scratch) return AETHER_E_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  char* p = (char*)scratch;
  Scratch s;
  s.cnt = (int32_t*)p; p += align_up((size_t)(Nn + 1) * 4, 256);
  s.fill = (int32_t*)p; p += align_up((size_t)(Nn + 1) * 4, 256);
  s.tmp = (int32_t*)p;
  AE_CHECK(cudaMemsetAsync(status, 0, 4, st));
  AE_CHECK(csr_by_key<int64_t>(recv, E, Nn, g->rowptr, g->perm, status, s, st));
  if (E > 0) {
    gather_kernel<int64_t><<<nblk(E, 256), 256, 0, st>>>(E, g->perm, send, recv, g->send, g->recv);
    AE_LAUNCH_CHECK();
  }
  AE_CHECK(csr_by_key<int32_t>(g->send, E, Nn, g->colptr, g->sperm, status, s, st));
  return 0;
}

extern "C" int aether_fc_edges(int64_t n, int64_t B, int64_t* send, int64_t* recv, void* stream) {
  if (n < 0 || B < 0) return AETHER_E_BADARG;
  const int64_t tot = n * (n - 1) * B;
  if (tot <= 0) return 0;
  fc_edges_kernel<<<nblk(tot, 256), 256, 0, (cudaStream_t)stream>>>(n, B, send, recv);
  AE_LAUNCH_CHECK();
  return 0;
}

extern "C" int aether_knn_edges(const float* pos, int32_t pos_stride, const int32_t* node_off,
                                const int64_t* edge_off, int32_t n_scenes, int32_t k, int64_t* send, int64_t* recv,
                                void* stream) {
  if (n_scenes <= 0) return 0;
  if (k <= 0 || pos_stride < 2) return AETHER_E_BADARG;
  // shared memory sized for the largest supported scene (1024 agents)
  knn_kernel<<<n_scenes, 64, 1024 * 2 * sizeof(float), (cudaStream_t)stream>>>(pos, pos_stride, node_off, edge_off, k,
                                                                               send, recv);
  AE_LAUNCH_CHECK();
  return 0;
}
