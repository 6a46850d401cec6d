// Backward of the localizer geometry (nn/state2state/aether.py:33-100, nn/utils/geometry.py:7-105,
// nn/utils/local_to_global.py:12-13) -- hand-derived reverse mode of geom.cu.
//
// Conventions matched to ATen's autograd formulas: atan2 and norm have zero gradient at the
// origin; clamp passes the gradient on the closed interval.  Where ATen would produce inf/NaN
// (acos/asin at exactly +-1) this code contributes 0 instead.
#include "geom.cuh"

namespace aether {

namespace {

__host__ __device__ __forceinline__ void atan2_bwd(float g, float y, float x, float& gy, float& gx) {
  const float den = x * x + y * y;
  if (den > 0.0f) { gy += g * x / den; gx += g * (-y) / den; }
}

// phi = acos(clamp(z / (|p| + eps), -1, 1)); adds d phi / d p * g into gp
template <int D>
__host__ __device__ __forceinline__ void phi_bwd(float g, const float* p, float* gp) {
  const float rho = norm2<D>(p);
  const float den = rho + AE_EPS_PHI;
  const float u = p[2] / den;
  if (u < -1.0f || u > 1.0f) return;
  const float om = 1.0f - u * u;
  if (!(om > 0.0f)) return;
  const float gu = -g / sqrtf(om);
  gp[2] += gu / den;
  if (rho > 0.0f) {
    const float grho = -gu * p[2] / (den * den);
#pragma unroll
    for (int k = 0; k < D; ++k) gp[k] += grho * p[k] / rho;
  }
}

template <int D>
__host__ __device__ inline void globalize_bwd_body(int64_t n, const float* R, const float* pred,
                                     const float* gout, float* gpred,
                                     float* gR) {
  float Rn[D * D], go[D], p[D], gp[D];
#pragma unroll
  for (int k = 0; k < D * D; ++k) Rn[k] = R[n * D * D + k];
#pragma unroll
  for (int k = 0; k < D; ++k) { go[k] = gout[n * D + k]; p[k] = pred[n * D + k]; }
  rot_t<D>(Rn, go, gp);
#pragma unroll
  for (int k = 0; k < D; ++k) gpred[n * D + k] = gp[k];
#pragma unroll
  for (int a = 0; a < D; ++a)
#pragma unroll
    for (int k = 0; k < D; ++k) gR[n * D * D + a * D + k] = go[a] * p[k];
}

// gsend[e] = [g_d D | g_vj D | g_fj D | dR_j D*D]   grecv[e] = [g_d D | g_a D | g_b D | dR_i D*D]
template <int D>
__host__ __device__ inline void edge_features_bwd_body(int64_t e, const int32_t* send, const int32_t* recv,
                                         const float* x, const float* vel,
                                         const float* f, const float* R,
                                         const float* gfeat, int ldf, float* gsend,
                                         float* grecv) {
  constexpr int O = D * (D - 1) / 2;
  constexpr int GS = 3 * D + D * D;
  const int64_t j = send[e], i = recv[e];
  const float* G = gfeat + (size_t)e * ldf;
  float Ri[D * D], Rj[D * D], d[D], vj[D], fj[D], dl[D], gdl[D], gd[D];
  float dRi[D * D], dRj[D * D];
#pragma unroll
  for (int k = 0; k < D * D; ++k) { Ri[k] = R[i * D * D + k]; Rj[k] = R[j * D * D + k]; dRi[k] = 0.f; dRj[k] = 0.f; }
#pragma unroll
  for (int k = 0; k < D; ++k) {
    d[k] = x[j * D + k] - x[i * D + k];
    vj[k] = vel[j * D + k];
    fj[k] = f[j * D + k];
  }
  rot_t<D>(Ri, d, dl);
  constexpr int c1 = D, c2 = D + O, c3 = D + O + 1, c4 = 2 * D + O, c5 = 3 * D + O, c7 = 5 * D + O, c8 = 6 * D + O;
#pragma unroll
  for (int a = 0; a < D; ++a) gdl[a] = G[a];
  atan2_bwd(G[c3], dl[1], dl[0], gdl[1], gdl[0]);
  if (D == 3) phi_bwd<D>(G[c3 + 1], dl, gdl);
  rot<D>(Ri, gdl, gd);
  const float dist = norm2<D>(d);
  if (dist > 0.0f) {
#pragma unroll
    for (int k = 0; k < D; ++k) gd[k] += G[c2] * d[k] / dist;
  }
  float gvj[D], gfj[D], gvl[D], gfl[D];
#pragma unroll
  for (int a = 0; a < D; ++a) { gvl[a] = G[c4 + a]; gfl[a] = G[c5 + a]; }
  rot<D>(Ri, gvl, gvj);
  rot<D>(Ri, gfl, gfj);
#pragma unroll
  for (int k = 0; k < D; ++k)
#pragma unroll
    for (int a = 0; a < D; ++a) dRi[k * D + a] += d[k] * gdl[a] + vj[k] * gvl[a] + fj[k] * gfl[a];
  // euler angles of M = R_i^T R_j
  float M[D * D], gM[D * D];
#pragma unroll
  for (int a = 0; a < D; ++a)
#pragma unroll
    for (int b = 0; b < D; ++b) {
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < D; ++k) s += Ri[k * D + a] * Rj[k * D + b];
      M[a * D + b] = s;
      gM[a * D + b] = 0.f;
    }
  atan2_bwd(G[c1] / AE_PI_F, M[1 * D + 0], M[0], gM[1 * D + 0], gM[0]);
  if (D == 3) {
    const float m20 = M[2 * D + 0];
    const float om = 1.0f - m20 * m20;
    if (om > 0.0f) gM[2 * D + 0] += -(G[c1 + 1] / AE_PI_F) / sqrtf(om);
    atan2_bwd(G[c1 + 2] / AE_PI_F, M[2 * D + 1], M[2 * D + 2], gM[2 * D + 1], gM[2 * D + 2]);
  }
#pragma unroll
  for (int k = 0; k < D; ++k)
#pragma unroll
    for (int a = 0; a < D; ++a) {
      float si = 0.f, sj = 0.f;
#pragma unroll
      for (int b = 0; b < D; ++b) {
        si += Rj[k * D + b] * gM[a * D + b];     // dR_i[k][a]
        sj += Ri[k * D + b] * gM[b * D + a];     // dR_j[k][a] = sum_b R_i[k][b] gM[b][a]
      }
      dRi[k * D + a] += si;
      dRj[k * D + a] += sj;
    }
  float* os = gsend + (size_t)e * GS;
  float* orr = grecv + (size_t)e * GS;
#pragma unroll
  for (int k = 0; k < D; ++k) {
    os[k] = gd[k]; os[D + k] = gvj[k]; os[2 * D + k] = gfj[k];
    orr[k] = gd[k]; orr[D + k] = G[c7 + k]; orr[2 * D + k] = G[c8 + k];
  }
#pragma unroll
  for (int k = 0; k < D * D; ++k) { os[3 * D + k] = dRj[k]; orr[3 * D + k] = dRi[k]; }
}

template <int D>
__host__ __device__ inline void node_frame_bwd_body(int64_t n, const float* vel, const float* f,
                                      const float* R, const float* gx0,
                                      const float* ssend, const float* srecv,
                                      const float* gRglob, const float* gout,
                                      float* gf, float* gx, float* gvel) {
  constexpr int GS = 3 * D + D * D;
  const float* S = ssend + (size_t)n * GS;
  const float* Rv = srecv + (size_t)n * GS;
  float v[D], ff[D], Rn[D * D], ga[D], gb[D], dR[D * D], gv[D], gff[D], t[D];
#pragma unroll
  for (int k = 0; k < D; ++k) {
    v[k] = vel[n * D + k];
    ff[k] = f[n * D + k];
    ga[k] = gx0[n * 3 * D + D + k] + Rv[D + k];
    gb[k] = gx0[n * 3 * D + 2 * D + k] + Rv[2 * D + k];
  }
#pragma unroll
  for (int k = 0; k < D * D; ++k) {
    Rn[k] = R[n * D * D + k];
    dR[k] = S[3 * D + k] + Rv[3 * D + k] + gRglob[n * D * D + k];
  }
  rot<D>(Rn, ga, gv);
  rot<D>(Rn, gb, t);
#pragma unroll
  for (int k = 0; k < D; ++k) {
    gv[k] += S[D + k];
    gff[k] = t[k] + S[2 * D + k];
#pragma unroll
    for (int c = 0; c < D; ++c) dR[k * D + c] += v[k] * ga[c] + ff[k] * gb[c];
  }
  // R(theta[,phi]) -> angles -> velocity
  float th = 0.f, ph = 0.f, Rtmp[D * D];
  heading_frame<D>(v, Rtmp, &th, &ph);
  const float ct = cosf(th), st = sinf(th);
  float gth;
  if (D == 2) {
    gth = dR[0] * (-st) + dR[1] * (-ct) + dR[2] * ct + dR[3] * (-st);
  } else {
    const float cp = cosf(ph), sp = sinf(ph);
    gth = dR[0] * (-cp * st) + dR[1] * (-ct) + dR[2] * (-sp * st) + dR[3] * (cp * ct) + dR[4] * (-st) + dR[5] * (sp * ct);
    const float gph = dR[0] * (-sp * ct) + dR[2] * (cp * ct) + dR[3] * (-sp * st) + dR[5] * (cp * st) + dR[6] * (-cp) +
                      dR[8] * (-sp);
    phi_bwd<D>(gph, v, gv);
  }
  atan2_bwd(gth, v[1], v[0], gv[1], gv[0]);
#pragma unroll
  for (int k = 0; k < D; ++k) {
    gf[n * D + k] = gff[k];
    gvel[n * D + k] = gv[k];
    gx[n * D + k] = gout[n * D + k] + S[k] - Rv[k];
  }
}

template <int D>
__global__ void add_field_input_grads_kernel(int64_t Nn, const float* __restrict__ gfin, float* gx, float* gvel) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Nn * D) return;
  const int64_t n = i / D;
  const int k = (int)(i % D);
  constexpr int W = 2 * D + 16;
  if (gx) gx[i] += gfin[n * W + k];
  if (gvel) gvel[i] += gfin[n * W + D + k];
}

template <int D>
__global__ void globalize_bwd_kernel(int64_t Nn, const float* __restrict__ R, const float* __restrict__ pred,
                                     const float* __restrict__ gout, float* __restrict__ gpred,
                                     float* __restrict__ gR) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= Nn) return;
  globalize_bwd_body<D>(n, R, pred, gout, gpred, gR);
}

template <int D>
__global__ void edge_features_bwd_kernel(int64_t E, const int32_t* __restrict__ send, const int32_t* __restrict__ recv,
                                         const float* __restrict__ x, const float* __restrict__ vel,
                                         const float* __restrict__ f, const float* __restrict__ R,
                                         const float* __restrict__ gfeat, int ldf, float* __restrict__ gsend,
                                         float* __restrict__ grecv) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  edge_features_bwd_body<D>(e, send, recv, x, vel, f, R, gfeat, ldf, gsend, grecv);
}

template <int D>
__global__ void node_frame_bwd_kernel(int64_t Nn, const float* __restrict__ vel, const float* __restrict__ f,
                                      const float* __restrict__ R, const float* __restrict__ gx0,
                                      const float* __restrict__ ssend, const float* __restrict__ srecv,
                                      const float* __restrict__ gRglob, const float* __restrict__ gout,
                                      float* __restrict__ gf, float* __restrict__ gx, float* __restrict__ gvel) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= Nn) return;
  node_frame_bwd_body<D>(n, vel, f, R, gx0, ssend, srecv, gRglob, gout, gf, gx, gvel);
}

inline unsigned nblk_gb(int64_t n, int b) { return (unsigned)((n + b - 1) / b); }

}  // namespace

int launch_globalize_bwd(int D, int64_t Nn, const float* R, const float* pred, const float* gout, float* gpred,
                         float* gR, cudaStream_t st) {
  if (Nn <= 0) return 0;
  if (D == 2) globalize_bwd_kernel<2><<<nblk_gb(Nn, 128), 128, 0, st>>>(Nn, R, pred, gout, gpred, gR);
  else globalize_bwd_kernel<3><<<nblk_gb(Nn, 128), 128, 0, st>>>(Nn, R, pred, gout, gpred, gR);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_edge_features_bwd(int D, int64_t E, const int32_t* send, const int32_t* recv, const float* x,
                             const float* vel, const float* f, const float* R, const float* gfeat, int ldf,
                             float* gsend, float* grecv, cudaStream_t st) {
  if (E <= 0) return 0;
  if (D == 2) edge_features_bwd_kernel<2><<<nblk_gb(E, 128), 128, 0, st>>>(E, send, recv, x, vel, f, R, gfeat, ldf, gsend, grecv);
  else edge_features_bwd_kernel<3><<<nblk_gb(E, 128), 128, 0, st>>>(E, send, recv, x, vel, f, R, gfeat, ldf, gsend, grecv);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_node_frame_bwd(int D, int64_t Nn, const float* vel, const float* f, const float* R, const float* gx0,
                          const float* ssend, const float* srecv, const float* gRglob, const float* gout, float* gf,
                          float* gx, float* gvel, cudaStream_t st) {
  if (Nn <= 0) return 0;
  if (D == 2) node_frame_bwd_kernel<2><<<nblk_gb(Nn, 128), 128, 0, st>>>(Nn, vel, f, R, gx0, ssend, srecv, gRglob, gout, gf, gx, gvel);
  else node_frame_bwd_kernel<3><<<nblk_gb(Nn, 128), 128, 0, st>>>(Nn, vel, f, R, gx0, ssend, srecv, gRglob, gout, gf, gx, gvel);
  AE_LAUNCH_CHECK();
  return 0;
}

int launch_add_field_input_grads(int D, int64_t Nn, const float* gfin, float* gx, float* gvel, cudaStream_t st) {
  if (Nn <= 0) return 0;
  if (D == 2) add_field_input_grads_kernel<2><<<nblk_gb(Nn * D, 256), 256, 0, st>>>(Nn, gfin, gx, gvel);
  else add_field_input_grads_kernel<3><<<nblk_gb(Nn * D, 256), 256, 0, st>>>(Nn, gfin, gx, gvel);
  AE_LAUNCH_CHECK();
  return 0;
}

}  // namespace aether
