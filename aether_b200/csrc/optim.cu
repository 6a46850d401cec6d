// Fused AdamW over all parameter tensors of the model in one launch (the reference's caller runs
// torch.optim.AdamW(lr=5e-4, weight_decay=1e-12), experiments/lorentz/main.py:164,292).
//
// Semantics are torch.optim.AdamW's (decoupled weight decay, bias-corrected moments, eps added after the
// square root):   p <- p (1 - lr wd);  m <- b1 m + (1 - b1) g;  v <- b2 v + (1 - b2) g^2
//                 p <- p - (lr / (1 - b1^t)) m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// The step count t lives on the device and is advanced by a one-thread launch in front of the update, so the pair can
// be captured in a CUDA graph (aether_b200/train.py) and replayed.
#include "common.cuh"

namespace aether {
namespace {

constexpr int kMaxTensors = 64;

struct AdamTable {
  float* p[kMaxTensors];
  const float* g[kMaxTensors];
  float* m[kMaxTensors];
  float* v[kMaxTensors];
  int64_t end[kMaxTensors];     // exclusive prefix end in units of 4 elements
  int64_t numel[kMaxTensors];
  int32_t n;
};

__global__ void adamw_tick_kernel(float* step) { *step += 1.0f; }

// omb1 = 1 - beta1, omb2 = 1 - beta2, decay = 1 - lr wd: formed in double on the host like torch does (1 - 0.999f in fp32
// is 4.7e-5 off)
__global__ void adamw_kernel(const __grid_constant__ AdamTable t, const float* __restrict__ step, double lr, double b1d, double b2d,
                             float eps, float omb1, float omb2, float decay) {
  const int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= t.end[t.n - 1]) return;
  int lo = 0, hi = t.n - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (u < t.end[mid]) hi = mid; else lo = mid + 1;
  }
  const int64_t i0 = (u - (lo ? t.end[lo - 1] : 0)) * 4;
  const float ts = *step;
  const double bc1 = 1.0 - pow(b1d, (double)ts), bc2 = 1.0 - pow(b2d, (double)ts);
  const float step_size = (float)(lr / bc1), inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  const float b2 = (float)b2d;
  float* p = t.p[lo];
  const float* g = t.g[lo];
  float* m = t.m[lo];
  float* v = t.v[lo];
  const int64_t n = t.numel[lo];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int64_t i = i0 + k;
    if (i >= n) break;
    const float gi = g[i];
    float pi = p[i] * decay;
    const float mi = m[i] + omb1 * (gi - m[i]);               // torch: exp_avg.lerp_(grad, 1 - beta1)
    const float vi = b2 * v[i] + omb2 * gi * gi;
    const float denom = sqrtf(vi) * inv_sqrt_bc2 + eps;
    pi -= step_size * (mi / denom);
    p[i] = pi; m[i] = mi; v[i] = vi;
  }
}

}  // namespace
}  // namespace aether

using namespace aether;

extern "C" int aether_adamw_step(int32_t n_tensors, float* const* params, const float* const* grads, float* const* exp_avg,
                                 float* const* exp_avg_sq, const int64_t* numel, float* step_dev, double lr, double beta1,
                                 double beta2, double eps, double weight_decay, void* stream) {
  if (n_tensors <= 0) return 0;
  if (n_tensors > kMaxTensors || !params || !grads || !exp_avg || !exp_avg_sq || !numel || !step_dev) return AETHER_E_BADARG;
  AdamTable t;
  int64_t tot = 0;
  int n = 0;
  for (int i = 0; i < n_tensors; ++i) {
    if (numel[i] <= 0) continue;
    if (!params[i] || !grads[i] || !exp_avg[i] || !exp_avg_sq[i]) return AETHER_E_BADARG;
    t.p[n] = params[i]; t.g[n] = grads[i]; t.m[n] = exp_avg[i]; t.v[n] = exp_avg_sq[i]; t.numel[n] = numel[i];
    tot += (numel[i] + 3) / 4;
    t.end[n] = tot;
    ++n;
  }
  for (int i = n; i < kMaxTensors; ++i) { t.p[i] = nullptr; t.g[i] = nullptr; t.m[i] = nullptr; t.v[i] = nullptr; t.end[i] = tot; t.numel[i] = 0; }
  t.n = n;
  cudaStream_t st = (cudaStream_t)stream;
  adamw_tick_kernel<<<1, 1, 0, st>>>(step_dev);
  AE_LAUNCH_CHECK();
  if (n == 0) return 0;
  // hyper-parameters arrive as doubles (Python floats): 1 - beta, 1 - lr wd and the bias corrections are formed in double
  // and rounded once, like torch.optim.AdamW does
  adamw_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(t, step_dev, lr, beta1, beta2, (float)eps, (float)(1.0 - beta1),
                                                                (float)(1.0 - beta2), (float)(1.0 - lr * weight_decay));
  AE_LAUNCH_CHECK();
  return 0;
}
