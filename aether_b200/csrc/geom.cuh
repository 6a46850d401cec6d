// Device-side geometry of the Aether localizer.  Operation order follows the reference so that
// fp32 results agree to rounding: nn/utils/geometry.py:7-105, nn/state2state/aether.py:33-100.
// This header is only included by files compiled with -fmad=false (no FMA contraction), which
// keeps products and sums rounded separately like the reference's eager ATen ops.
#pragma once
#include "common.cuh"

namespace aether {

#define AE_PI_F 3.14159265358979323846f
#define AE_2PI_F 6.283185307179586f
#define AE_EPS_PHI 1e-7f   /* nn/utils/geometry.py:63 */

// y = R^T x  (einsum('ij,j->i', R.transpose, x): y_a = sum_k R[k][a] x[k], k ascending)
template <int D>
__host__ __device__ __forceinline__ void rot_t(const float* R, const float* x, float* y) {
#pragma unroll
  for (int a = 0; a < D; ++a) {
    float s = 0.0f;   // accumulate from +0 like the reference's bmm (keeps the sign of zero sums)
#pragma unroll
    for (int k = 0; k < D; ++k) s = s + R[k * D + a] * x[k];
    y[a] = s;
  }
}
// y = R x
template <int D>
__host__ __device__ __forceinline__ void rot(const float* R, const float* x, float* y) {
#pragma unroll
  for (int a = 0; a < D; ++a) {
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < D; ++k) s = s + R[a * D + k] * x[k];
    y[a] = s;
  }
}

template <int D>
__host__ __device__ __forceinline__ float norm2(const float* v) {
  float s = v[0] * v[0];
#pragma unroll
  for (int k = 1; k < D; ++k) s = s + v[k] * v[k];
  return sqrtf(s);
}

__host__ __device__ __forceinline__ float clamp1(float z) { return fminf(fmaxf(z, -1.0f), 1.0f); }

// Heading frame of a velocity.  geometry.py:69-73 -> :37-66 (theta wrapped to [0,2pi)) -> :7-34.
// Also returns the angles (needed by the backward pass).
template <int D>
__host__ __device__ __forceinline__ void heading_frame(const float* v, float* R, float* theta_out, float* phi_out) {
  float th = atan2f(v[1], v[0]);
  if (th < 0.0f) th = th + AE_2PI_F;
  const float ct = cosf(th), st = sinf(th);
  if (theta_out) *theta_out = th;
  if (D == 2) {
    R[0] = ct; R[1] = -st;
    R[2] = st; R[3] = ct;
  } else {
    const float rho = norm2<D>(v);
    const float ph = acosf(clamp1(v[2] / (rho + AE_EPS_PHI)));
    if (phi_out) *phi_out = ph;
    const float cp = cosf(ph), sp = sinf(ph);
    R[0] = cp * ct; R[1] = -st; R[2] = sp * ct;
    R[3] = cp * st; R[4] = ct;  R[5] = sp * st;
    R[6] = -sp;     R[7] = 0.f; R[8] = cp;
  }
}

// ---- per-node / per-edge bodies shared by the geometry kernels (geom.cu) and the single-kernel rollout (small.cu) ----

// R[n], x0[n] = [0_D, R^T v, R^T f]   nn/state2state/aether.py:33-50
template <int D>
__host__ __device__ inline void node_frame_body(int64_t n, const float* vel, const float* f,
                                  float* R, float* x0) {
  float v[D], ff[D], Rn[D * D], a[D], b[D];
#pragma unroll
  for (int k = 0; k < D; ++k) { v[k] = vel[n * D + k]; ff[k] = f[n * D + k]; }
  heading_frame<D>(v, Rn, nullptr, nullptr);
  rot_t<D>(Rn, v, a);
  rot_t<D>(Rn, ff, b);
#pragma unroll
  for (int k = 0; k < D * D; ++k) R[n * D * D + k] = Rn[k];
#pragma unroll
  for (int k = 0; k < D; ++k) {
    x0[n * 3 * D + k] = 0.0f;
    x0[n * 3 * D + D + k] = a[k];
    x0[n * 3 * D + 2 * D + k] = b[k];
  }
}

// One thread per receiver-sorted edge.  Column order (aether.py:81-99,177):
// [R_i^T(p_j-p_i) D | euler(R_i^T R_j)/pi O | |p_j-p_i| 1 | theta(,phi) D-1 | R_i^T v_j D | R_i^T f_j D |
//  rel_feat[i] 3D | edge_attr_orig 2 | zero pad]
template <int D>
__host__ __device__ inline void edge_features_core(int64_t j, int64_t i, const float* x, const float* vel, const float* f,
                                                   const float* R, const float* x0, float orig0, float orig1,
                                                   float* out /* [(FE+3)/4*4], tail zeroed */) {
  constexpr int O = D * (D - 1) / 2;
  constexpr int FE = 7 * D + O + 2;
  constexpr int LD = (FE + 3) / 4 * 4;
  float Ri[D * D], Rj[D * D], d[D], vj[D], fj[D];
#pragma unroll
  for (int k = 0; k < D * D; ++k) { Ri[k] = R[i * D * D + k]; Rj[k] = R[j * D * D + k]; }
#pragma unroll
  for (int k = 0; k < D; ++k) {
    d[k] = x[j * D + k] - x[i * D + k];
    vj[k] = vel[j * D + k];
    fj[k] = f[j * D + k];
  }
  int c = 0;
  float dl[D];
  rot_t<D>(Ri, d, dl);
#pragma unroll
  for (int k = 0; k < D; ++k) out[c++] = dl[k];
  // M = R_i^T R_j ; euler (geometry.py:76-101)
  if (D == 2) {
    const float m00 = Ri[0] * Rj[0] + Ri[2] * Rj[2];
    const float m10 = Ri[1] * Rj[0] + Ri[3] * Rj[2];
    out[c++] = atan2f(m10, m00) / AE_PI_F;
  } else {
    auto Mab = [&](int a, int b) { return Ri[0 * D + a] * Rj[0 * D + b] + Ri[1 * D + a] * Rj[1 * D + b] +
                                          Ri[2 * D + a] * Rj[2 * D + b]; };
    const float m00 = Mab(0, 0), m10 = Mab(1, 0), m20 = Mab(2, 0), m21 = Mab(2, 1), m22 = Mab(2, 2);
    out[c++] = atan2f(m10, m00) / AE_PI_F;
    out[c++] = asinf(-m20) / AE_PI_F;
    out[c++] = atan2f(m21, m22) / AE_PI_F;
  }
  out[c++] = norm2<D>(d);
  out[c++] = atan2f(dl[1], dl[0]);                       // symmetric theta (aether.py:73)
  if (D == 3) {
    const float rho = norm2<D>(dl);
    out[c++] = acosf(clamp1(dl[2] / (rho + AE_EPS_PHI)));
  }
  float t[D];
  rot_t<D>(Ri, vj, t);
#pragma unroll
  for (int k = 0; k < D; ++k) out[c++] = t[k];
  rot_t<D>(Ri, fj, t);
#pragma unroll
  for (int k = 0; k < D; ++k) out[c++] = t[k];
#pragma unroll
  for (int k = 0; k < 3 * D; ++k) out[c++] = x0[i * 3 * D + k];
  out[c++] = orig0;
  out[c++] = orig1;
#pragma unroll
  for (int k = FE; k < LD; ++k) out[k] = 0.0f;
}

template <int D>
__host__ __device__ inline void edge_features_values(int64_t e, const int32_t* send, const int32_t* recv,
                                     const int32_t* perm, const float* x,
                                     const float* vel, const float* f,
                                     const float* R, const float* x0,
                                     const float* orig, float* out /* [(FE+3)/4*4], tail zeroed */) {
  const int64_t pe = perm[e];
  edge_features_core<D>(send[e], recv[e], x, vel, f, R, x0, orig[pe * 2 + 0], orig[pe * 2 + 1], out);
}

template <int D>
__host__ __device__ inline void edge_features_body(int64_t e, const int32_t* send, const int32_t* recv,
                                     const int32_t* perm, const float* x,
                                     const float* vel, const float* f,
                                     const float* R, const float* x0,
                                     const float* orig, float* feat, int ldf) {
  constexpr int LD = (7 * D + D * (D - 1) / 2 + 2 + 3) / 4 * 4;
  float out[LD];
  edge_features_values<D>(e, send, recv, perm, x, vel, f, R, x0, orig, out);
  float4* dst = reinterpret_cast<float4*>(feat + (size_t)e * ldf);
#pragma unroll
  for (int k = 0; k < LD / 4; ++k) dst[k] = make_float4(out[4 * k], out[4 * k + 1], out[4 * k + 2], out[4 * k + 3]);
}

// out = x + R pred   (local_to_global.py:12-13, aether.py:185)
template <int D>
__host__ __device__ inline void globalize_body(int64_t n, const float* x, const float* R,
                                 const float* pred, float* out) {
  float Rn[D * D], p[D], g[D];
#pragma unroll
  for (int k = 0; k < D * D; ++k) Rn[k] = R[n * D * D + k];
#pragma unroll
  for (int k = 0; k < D; ++k) p[k] = pred[n * D + k];
  rot<D>(Rn, p, g);
#pragma unroll
  for (int k = 0; k < D; ++k) out[n * D + k] = x[n * D + k] + g[k];
}

}  // namespace aether
