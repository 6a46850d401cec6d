// Throughput of the legacy tensor-core path (mma.sync.m16n8k16 bf16, fp32 accumulate) and of FFMA, registers only.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu ; run on the GPU box.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void mma_loop(float* out, int iters) {
  float d[8][4];
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) d[i][j] = 0.f;
  uint32_t a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(d[i][0]), "+f"(d[i][1]), "+f"(d[i][2]), "+f"(d[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float s = 0.f;
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += d[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void ffma_loop(float* out, int iters) {
  float d[32];
  for (int i = 0; i < 32; ++i) d[i] = threadIdx.x * 1e-3f + i;
  const float a = 1.0001f, b = 0.5f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 32; ++i) d[i] = fmaf(d[i], a, b);
  }
  float s = 0.f;
  for (int i = 0; i < 32; ++i) s += d[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * sizeof(float));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int warps = 4; warps <= 16; warps *= 2) {
    const int iters = 20000, blocks = 148 * (warps > 8 ? 2 : 1), threads = 32 * (warps > 8 ? warps / 2 : warps);
    float ms;
    mma_loop<<<blocks, threads>>>(out, 100); cudaDeviceSynchronize();
    cudaEventRecord(e0); mma_loop<<<blocks, threads>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    double flop = 2.0 * 16 * 8 * 16 * 8.0 * iters * (double)blocks * threads / 32;
    printf("mma.sync m16n8k16 bf16: %2d warps/SM: %.1f TFLOP/s (%.3f ms)\n", warps, flop / ms * 1e-9, ms);
    ffma_loop<<<blocks, threads>>>(out, 100); cudaDeviceSynchronize();
    cudaEventRecord(e0); ffma_loop<<<blocks, threads>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    flop = 2.0 * 32.0 * iters * (double)blocks * threads;
    printf("FFMA                  : %2d warps/SM: %.1f TFLOP/s (%.3f ms)\n", warps, flop / ms * 1e-9, ms);
  }
  return 0;
}
