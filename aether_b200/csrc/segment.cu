// Segment kernels on the receiver-sorted edge list: atomics-free mean / sum over CSR segments
// (replaces torch_scatter.scatter(reduce="mean"), nn/state2state/locs/locs.py:236-238, and its
// backward), plus the small elementwise backward helpers.
#include "common.cuh"

namespace aether {

namespace {

// One warp per node; lanes stride the feature dimension with float4 (coalesced 512 B per row step).
// Rows of a segment are contiguous (receiver-sorted), summed in edge order = the order
// scatter_add_ visits them on the CPU, then divided by the clamped in-degree.
template <bool SILU, bool MEAN>
__global__ void segreduce_kernel(int64_t Nn, int W, const int32_t* __restrict__ ptr, const int32_t* __restrict__ list,
                                 const float* __restrict__ src, int64_t lds, const float* base,
                                 int64_t ldb, float* y, int64_t ldy, int accumulate) {
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (n >= Nn) return;
  const int b = ptr[n], e = ptr[n + 1];
  const bool vec = (W % 4 == 0) && (lds % 4 == 0);
  if (vec) {
    for (int c = lane * 4; c < W; c += 128) {
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = b; k < e; ++k) {
        const int64_t row = list ? list[k] : k;
        const float4 v = *reinterpret_cast<const float4*>(src + (size_t)row * lds + c);
        if (SILU) { s.x += silu_f(v.x); s.y += silu_f(v.y); s.z += silu_f(v.z); s.w += silu_f(v.w); }
        else { s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
      }
      float r[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float v = r[q];
        if (MEAN) { const int deg = e - b; v = v / (float)(deg < 1 ? 1 : deg); }
        if (base) v = base[(size_t)n * ldb + c + q] + v;
        float* o = y + (size_t)n * ldy + c + q;
        *o = accumulate ? *o + v : v;
      }
    }
  } else {
    for (int c = lane; c < W; c += 32) {
      float s = 0.f;
      for (int k = b; k < e; ++k) {
        const int64_t row = list ? list[k] : k;
        const float v = src[(size_t)row * lds + c];
        s += SILU ? silu_f(v) : v;
      }
      if (MEAN) { const int deg = e - b; s = s / (float)(deg < 1 ? 1 : deg); }
      if (base) s = base[(size_t)n * ldb + c] + s;
      float* o = y + (size_t)n * ldy + c;
      *o = accumulate ? *o + s : s;
    }
  }
}

__global__ void edge_dpre2_kernel(int64_t E, int H, const int32_t* __restrict__ recv,
                                  const int32_t* __restrict__ rowptr, const float* ge,
                                  const float* __restrict__ gy, const float* __restrict__ pre2, float* g) {
  const int hq = H / 4;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= E * hq) return;
  const int64_t k = i / hq;
  const int c = (int)(i % hq) * 4;
  const int r = recv[k];
  const int deg = rowptr[r + 1] - rowptr[r];
  const float dg = (float)(deg < 1 ? 1 : deg);
  const float4 a = *reinterpret_cast<const float4*>(gy + (size_t)r * H + c);
  const float4 z = *reinterpret_cast<const float4*>(pre2 + (size_t)k * H + c);
  float4 o = make_float4(a.x / dg, a.y / dg, a.z / dg, a.w / dg);
  if (ge) {
    const float4 b = *reinterpret_cast<const float4*>(ge + (size_t)k * H + c);
    o.x += b.x; o.y += b.y; o.z += b.z; o.w += b.w;
  }
  o.x *= dsilu_f(z.x); o.y *= dsilu_f(z.y); o.z *= dsilu_f(z.z); o.w *= dsilu_f(z.w);
  *reinterpret_cast<float4*>(g + (size_t)k * H + c) = o;
}

// gemb[r, c] = sum over nodes with class r of gfin[n, 2D + c].  Each block of 1024 threads takes a contiguous
// slab of nodes: thread (lane, c) walks nodes lane, lane+64, ... of the slab keeping one sum per class; lanes,
// then blocks, are combined in a fixed order (deterministic, no atomics).
__global__ void embedding_bwd_kernel(int D, int64_t Nn, int64_t slab, const float* __restrict__ charges,
                                     const float* __restrict__ gfin, float* __restrict__ partial) {
  __shared__ float part[64][49];
  const int c = threadIdx.x & 15, ln = threadIdx.x >> 4;   // 64 node lanes x 16 columns
  const int W = 2 * D + 16;
  const int64_t n0 = (int64_t)blockIdx.x * slab, n1 = n0 + slab < Nn ? n0 + slab : Nn;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f;
  for (int64_t n = n0 + ln; n < n1; n += 64) {
    long long idx = (long long)(charges[n] + 1.0f);
    idx = idx < 0 ? 0 : (idx > 2 ? 2 : idx);
    const float v = gfin[(size_t)n * W + 2 * D + c];
    if (idx == 0) s0 += v; else if (idx == 1) s1 += v; else s2 += v;
  }
  part[ln][c] = s0; part[ln][16 + c] = s1; part[ln][32 + c] = s2;
  __syncthreads();
  if (threadIdx.x < 48) {
    float a = 0.f;
    for (int i = 0; i < 64; ++i) a += part[i][threadIdx.x];
    partial[(size_t)blockIdx.x * 48 + threadIdx.x] = a;
  }
}
__global__ void embedding_bwd_reduce_kernel(const float* __restrict__ partial, int blocks, float* __restrict__ gemb) {
  if (threadIdx.x < 48) {
    float a = 0.f;
    for (int b = 0; b < blocks; ++b) a += partial[(size_t)b * 48 + threadIdx.x];
    gemb[threadIdx.x] = a;
  }
}

__global__ void fill_kernel(float* p, size_t n, float v) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

}  // namespace

int launch_segmean_silu(int64_t Nn, int H, const int32_t* rowptr, const float* src, int64_t lds, const float* base,
                        int64_t ldb, float* y, int64_t ldy, cudaStream_t st) {
  if (Nn <= 0) return 0;
  const unsigned blocks = (unsigned)((Nn * 32 + 255) / 256);
  segreduce_kernel<true, true><<<blocks, 256, 0, st>>>(Nn, H, rowptr, nullptr, src, lds, base, ldb, y, ldy, 0);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_segsum(int64_t Nn, int W, const int32_t* ptr, const int32_t* list, const float* src, int64_t lds,
                  float* out, int64_t ldo, int accumulate, cudaStream_t st) {
  if (Nn <= 0) return 0;
  const unsigned blocks = (unsigned)((Nn * 32 + 255) / 256);
  segreduce_kernel<false, false><<<blocks, 256, 0, st>>>(Nn, W, ptr, list, src, lds, nullptr, 0, out, ldo, accumulate);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_edge_dpre2(int64_t E, int H, const int32_t* recv, const int32_t* rowptr, const float* ge, const float* gy,
                      const float* pre2, float* g, cudaStream_t st) {
  if (E <= 0) return 0;
  if (H % 4) return AETHER_E_BADARG;
  const int64_t tot = E * (H / 4);
  edge_dpre2_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(E, H, recv, rowptr, ge, gy, pre2, g);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_embedding_bwd(int D, int64_t Nn, const float* charges, const float* gfin, float* gemb, float* partial,
                         size_t partial_floats, cudaStream_t st) {
  int blocks = (int)((Nn + 4095) / 4096);
  if (blocks < 1) blocks = 1;
  if (blocks > 256) blocks = 256;
  if (partial_floats < (size_t)blocks * 48) return AETHER_E_WORKSPACE;
  const int64_t slab = (Nn + blocks - 1) / blocks;
  embedding_bwd_kernel<<<blocks, 1024, 0, st>>>(D, Nn, slab, charges, gfin, partial);
  AE_LAUNCH_CHECK();
  embedding_bwd_reduce_kernel<<<1, 64, 0, st>>>(partial, blocks, gemb);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_fill(float* p, size_t n, float v, cudaStream_t st) {
  if (n == 0) return 0;
  fill_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(p, n, v);
  AE_LAUNCH_CHECK();
  return 0;
}

}  // namespace aether
