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

// Sorted copies of the edge list.  Memory safety for malformed lists (the reference raises IndexError on them): edges
// whose receiver is out of range were skipped by count/place, so the tail of `perm` holds the zeros it was cleared to;
// out-of-range senders are clamped into [0, Nn) and flagged in the status word.  The step then computes garbage for
// those edges but never reads out of bounds; PreparedGraph.validate() (or AETHER_B200_VALIDATE=1) turns the status
// word into the IndexError.
template <typename T>
__global__ void gather_kernel(int64_t E, int64_t Nn, const int32_t* __restrict__ perm, const T* __restrict__ a,
                              const T* __restrict__ b, int32_t* __restrict__ oa, int32_t* __restrict__ ob,
                              int32_t* status) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  int32_t p = perm[e];
  if (p < 0 || p >= E) p = 0;
  long long va = (long long)a[p], vb = (long long)b[p];
  if (va < 0 || va >= Nn || vb < 0 || vb >= Nn) {
    atomicOr(status, 2);
    va = va < 0 ? 0 : (va >= Nn ? Nn - 1 : va);
    vb = vb < 0 ? 0 : (vb >= Nn ? Nn - 1 : vb);
    if (Nn <= 0) { va = 0; vb = 0; }
  }
  oa[e] = (int32_t)va;
  ob[e] = (int32_t)vb;
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
        const float d = knn_dist(xi, yi, sp[2 * j], sp[2 * j + 1]);
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

// ---- radius graph capped at the k nearest, for scenes too large for the brute-force kernel above --------------
// Uniform cell list with cell edge >= radius: every neighbour within the radius lies in the 3^D cells around the
// query's cell.  Squared distances are accumulated left to right with separate roundings (no FMA), so they are
// bit-identical to the numpy restatement in oracle/; candidates are ranked by (d2, index), which makes the result
// independent of the (atomic) order in which particles were appended to their cells.
struct CellGrid { float lo[3]; float inv; int dim[3]; int D; };

__device__ __forceinline__ int cell_coord(float p, float lo, float inv, int dim) {
  int c = (int)floorf(__fmul_rn(__fsub_rn(p, lo), inv));
  return c < 0 ? 0 : (c >= dim ? dim - 1 : c);
}
__device__ __forceinline__ int cell_of(const float* __restrict__ p, const CellGrid& g) {
  int c = cell_coord(p[0], g.lo[0], g.inv, g.dim[0]) + g.dim[0] * cell_coord(p[1], g.lo[1], g.inv, g.dim[1]);
  if (g.D == 3) c += g.dim[0] * g.dim[1] * cell_coord(p[2], g.lo[2], g.inv, g.dim[2]);
  return c;
}

__global__ void cell_count_kernel(const float* __restrict__ pos, int stride, int64_t n, CellGrid g, int32_t* cell_cnt,
                                  int32_t* __restrict__ cell_id) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = cell_of(pos + (size_t)i * stride, g);
  cell_id[i] = c;
  atomicAdd(&cell_cnt[c], 1);
}

__global__ void cell_fill_kernel(int64_t n, const int32_t* __restrict__ cell_id, const int32_t* __restrict__ cell_ptr,
                                 int32_t* fill, int32_t* __restrict__ cell_items) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = cell_id[i];
  cell_items[cell_ptr[c] + atomicAdd(&fill[c], 1)] = (int32_t)i;
}

__device__ __forceinline__ bool cand_less(float da, int ja, float db, int jb) { return da < db || (da == db && ja < jb); }

// One warp per query particle; lane l holds the l-th best (d2, index) found so far.
__global__ void radius_knn_kernel(const float* __restrict__ pos, int stride, int64_t n, CellGrid g, float r2, int k,
                                  const int32_t* __restrict__ cell_ptr, const int32_t* __restrict__ cell_items,
                                  int32_t* __restrict__ nbr, int32_t* __restrict__ deg) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (i >= n) return;
  const float* pi = pos + (size_t)i * stride;
  const float xi = pi[0], yi = pi[1], zi = g.D == 3 ? pi[2] : 0.0f;
  const int cx = cell_coord(xi, g.lo[0], g.inv, g.dim[0]);
  const int cy = cell_coord(yi, g.lo[1], g.inv, g.dim[1]);
  const int cz = g.D == 3 ? cell_coord(zi, g.lo[2], g.inv, g.dim[2]) : 0;
  float bd = INFINITY;
  int bj = 0x7fffffff;
  const int zlo = g.D == 3 ? cz - 1 : 0, zhi = g.D == 3 ? cz + 1 : 0;
  for (int z = zlo; z <= zhi; ++z) {
    if (g.D == 3 && (z < 0 || z >= g.dim[2])) continue;
    for (int y = cy - 1; y <= cy + 1; ++y) {
      if (y < 0 || y >= g.dim[1]) continue;
      // the three x-neighbours are consecutive cells: one contiguous item range
      const int x0 = cx > 0 ? cx - 1 : 0, x1 = cx + 1 < g.dim[0] ? cx + 1 : g.dim[0] - 1;
      const int cbase = (z * g.dim[1] + y) * g.dim[0];
      const int beg = cell_ptr[cbase + x0], end = cell_ptr[cbase + x1 + 1];
      for (int b = beg; b < end; b += 32) {
        const int idx = b + lane;
        float dc = INFINITY;
        int jc = 0x7fffffff;
        if (idx < end) {
          jc = cell_items[idx];
          const float* pj = pos + (size_t)jc * stride;
          const float dx = __fsub_rn(xi, pj[0]), dy = __fsub_rn(yi, pj[1]);
          float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
          if (g.D == 3) { const float dz = __fsub_rn(zi, pj[2]); d2 = __fadd_rn(d2, __fmul_rn(dz, dz)); }
          dc = (jc != (int)i && d2 <= r2) ? d2 : INFINITY;
        }
        // insert, one at a time, every candidate that beats the current k-th best
        for (;;) {
          const float wd = __shfl_sync(0xffffffffu, bd, k - 1);
          const int wj = __shfl_sync(0xffffffffu, bj, k - 1);
          const unsigned m = __ballot_sync(0xffffffffu, dc < INFINITY && cand_less(dc, jc, wd, wj));
          if (!m) break;
          const int src = __ffs(m) - 1;
          const float dn = __shfl_sync(0xffffffffu, dc, src);
          const int jn = __shfl_sync(0xffffffffu, jc, src);
          if (lane == src) dc = INFINITY;
          const float du = __shfl_up_sync(0xffffffffu, bd, 1);
          const int ju = __shfl_up_sync(0xffffffffu, bj, 1);
          if (cand_less(dn, jn, bd, bj)) {
            if (lane > 0 && cand_less(dn, jn, du, ju)) { bd = du; bj = ju; }
            else { bd = dn; bj = jn; }
          }
        }
      }
    }
  }
  const unsigned have = __ballot_sync(0xffffffffu, lane < k && bd < INFINITY);
  if (lane < k) nbr[(size_t)i * k + lane] = bd < INFINITY ? bj : -1;
  if (lane == 0) deg[i] = __popc(have);
}

__global__ void radius_fill_kernel(const int32_t* __restrict__ nbr, const int32_t* __restrict__ off, int64_t n, int k,
                                   int64_t* __restrict__ send, int64_t* __restrict__ recv) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * k) return;
  const int64_t i = t / k;
  const int j = (int)(t % k);
  const int32_t o = off[i];
  if (j < off[i + 1] - o) { send[o + j] = i; recv[o + j] = nbr[t]; }
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
  if (E > 0) {
    // defined contents for the slots that edges with an out-of-range key never fill (see gather_kernel)
    AE_CHECK(cudaMemsetAsync(g->perm, 0, (size_t)E * 4, st));
    AE_CHECK(cudaMemsetAsync(g->sperm, 0, (size_t)E * 4, st));
  }
  AE_CHECK(csr_by_key<int64_t>(recv, E, Nn, g->rowptr, g->perm, status, s, st));
  if (E > 0) {
    gather_kernel<int64_t><<<nblk(E, 256), 256, 0, st>>>(E, Nn, g->perm, send, recv, g->send, g->recv, status);
    AE_LAUNCH_CHECK();
  }
  AE_CHECK(csr_by_key<int32_t>(g->send, E, Nn, g->colptr, g->sperm, status, s, st));
  return 0;
}

// ---- aether_graph_prepare without any allocation by anybody: the scratch is a file-scope device array ----------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  The array is sized for the stated
// maximum and is held by every GPU that loads the library until the process exits (2.5 MB).  In host code the array's
// NAME is a host-side stand-in, not a device address: the address is asked for once per device with
// cudaGetSymbolAddress (that device current) and only that pointer is handed to kernels.
namespace aether {
namespace {
constexpr size_t static_scratch_bytes(int64_t Nn, int64_t E) {
  return ((size_t)(Nn + 1) * 4 + 255) / 256 * 256 * 2 + ((size_t)E * 4 + 255) / 256 * 256;
}
constexpr size_t kStaticScratchBytes = static_scratch_bytes(AETHER_GRAPH_STATIC_MAX_NODES, AETHER_GRAPH_STATIC_MAX_EDGES);
static __device__ __align__(256) unsigned char g_graph_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
std::atomic<void*> g_static_scratch_ptr[kMaxDevices];     // device address of the array on device i; null: not asked yet

// Device address of the array on the CURRENT device (looked up once per device; no allocation: the array is part of the
// loaded module).
int static_scratch(void** out) {
  int dev = -1;
  AE_CHECK((int)cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return AETHER_E_UNSUPPORTED;
  void* p = g_static_scratch_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    AE_CHECK((int)cudaGetSymbolAddress(&p, g_graph_static_scratch));
    if (!p) return AETHER_E_WORKSPACE;
    g_static_scratch_ptr[dev].store(p, std::memory_order_release);   // racing threads store the same value
  }
  *out = p;
  return 0;
}
}  // namespace
}  // namespace aether

extern "C" int aether_graph_prepare_static(const int64_t* send, const int64_t* recv, int64_t E, int64_t Nn, AetherGraph* g,
                                           int32_t* status, void* stream) {
  if (E < 0 || Nn < 0) return AETHER_E_BADARG;
  if (Nn > AETHER_GRAPH_STATIC_MAX_NODES || E > AETHER_GRAPH_STATIC_MAX_EDGES) return AETHER_E_TOO_LARGE;
  void* scratch = nullptr;
  AE_CHECK(static_scratch(&scratch));
  // the existing entry point, unchanged, on the static array: it clears the counters it needs at zero itself
  // (csr_by_key) and overwrites the rest before reading it, so the previous call's contents do not matter
  return aether_graph_prepare(send, recv, E, Nn, g, status, scratch, kStaticScratchBytes, stream);
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

// ---- radius graph (large single scenes, BASELINE config 5) ----------------------------------------------------
namespace {
struct RadiusScratch { int32_t *cell_cnt, *cell_ptr, *fill, *cell_id, *items, *deg; size_t bytes; };
RadiusScratch carve_radius(int64_t n, int64_t cells, void* base) {
  RadiusScratch r;
  char* p = (char*)base;
  size_t off = 0;
  auto take = [&](size_t count) { int32_t* q = (int32_t*)(p + off); off += align_up(count * 4, 256); return q; };
  r.cell_cnt = take(cells + 1); r.cell_ptr = take(cells + 1); r.fill = take(cells + 1);
  r.cell_id = take(n > 0 ? n : 1); r.items = take(n > 0 ? n : 1); r.deg = take(n + 1);
  r.bytes = off;
  return r;
}
bool grid_ok(int32_t D, const int32_t* dim, int64_t* cells) {
  if (!dim || (D != 2 && D != 3)) return false;
  int64_t c = 1;
  for (int a = 0; a < D; ++a) { if (dim[a] < 1) return false; c *= dim[a]; if (c >= ((int64_t)1 << 30)) return false; }
  *cells = c;
  return true;
}
}  // namespace

extern "C" size_t aether_radius_graph_scratch_bytes(int64_t n, int32_t num_dims, const int32_t* grid_dim) {
  int64_t cells;
  if (n < 0 || !grid_ok(num_dims, grid_dim, &cells)) return 0;
  return carve_radius(n, cells, nullptr).bytes + 256;
}

extern "C" int aether_radius_graph_neighbors(const float* pos, int32_t pos_stride, int32_t num_dims, int64_t n,
                                             float radius, float cell_size, int32_t k, const float* box_lo,
                                             const int32_t* grid_dim, int32_t* nbr, int32_t* off, void* scratch, size_t scratch_bytes,
                                             void* stream) {
  int64_t cells;
  if (n < 0 || n >= ((int64_t)1 << 31) / (k > 0 ? k : 1) || k < 1 || k > 32 || !(radius > 0.0f) || !(cell_size >= radius) || !box_lo ||
      pos_stride < num_dims || !grid_ok(num_dims, grid_dim, &cells))
    return AETHER_E_BADARG;
  if (!scratch || ((uintptr_t)scratch & 255) || scratch_bytes < carve_radius(n, cells, nullptr).bytes) return AETHER_E_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  const RadiusScratch s = carve_radius(n, cells, scratch);
  CellGrid g;
  g.D = num_dims; g.inv = 1.0f / cell_size;
  for (int a = 0; a < 3; ++a) { g.lo[a] = a < num_dims ? box_lo[a] : 0.0f; g.dim[a] = a < num_dims ? grid_dim[a] : 1; }
  AE_CHECK(cudaMemsetAsync(s.cell_cnt, 0, (size_t)(cells + 1) * 4, st));
  AE_CHECK(cudaMemsetAsync(s.fill, 0, (size_t)(cells + 1) * 4, st));
  if (n > 0) {
    cell_count_kernel<<<nblk(n, 256), 256, 0, st>>>(pos, pos_stride, n, g, s.cell_cnt, s.cell_id);
    AE_LAUNCH_CHECK();
  }
  scan_kernel<<<1, 1024, 0, st>>>(s.cell_cnt, cells, s.cell_ptr);
  AE_LAUNCH_CHECK();
  if (n > 0) {
    cell_fill_kernel<<<nblk(n, 256), 256, 0, st>>>(n, s.cell_id, s.cell_ptr, s.fill, s.items);
    AE_LAUNCH_CHECK();
    radius_knn_kernel<<<nblk(n * 32, 256), 256, 0, st>>>(pos, pos_stride, n, g, radius * radius, k, s.cell_ptr, s.items,
                                                         nbr, s.deg);
    AE_LAUNCH_CHECK();
  }
  scan_kernel<<<1, 1024, 0, st>>>(s.deg, n, off);
  AE_LAUNCH_CHECK();
  return 0;
}

extern "C" int aether_radius_graph_fill(const int32_t* nbr, const int32_t* off, int64_t n, int32_t k, int64_t* send,
                                        int64_t* recv, void* stream) {
  if (n < 0 || k < 1 || k > 32) return AETHER_E_BADARG;
  if (n == 0) return 0;
  radius_fill_kernel<<<nblk(n * k, 256), 256, 0, (cudaStream_t)stream>>>(nbr, off, n, k, send, recv);
  AE_LAUNCH_CHECK();
  return 0;
}
