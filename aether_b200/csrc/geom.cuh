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

}  // namespace aether
