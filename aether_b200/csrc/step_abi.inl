// extern "C" surface of the step (included by step.cu).  See include/aether_b200.h.
using namespace aether;

extern "C" int aether_abi_version(void) { return AETHER_B200_ABI_VERSION; }

extern "C" const char* aether_error_string(int code) {
  switch (code) {
    case 0: return "success";
    case AETHER_E_BADARG: return "aether_b200: bad argument";
    case AETHER_E_WORKSPACE: return "aether_b200: workspace or scratch buffer too small";
    case AETHER_E_UNSUPPORTED: return "aether_b200: configuration not supported by this kernel";
    case AETHER_E_TOO_LARGE: return "aether_b200: size exceeds the maximum of this no-allocation entry point (static scratch: at most 65536 nodes and 524288 edges); use the variant that takes caller-provided scratch";
    default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "aether_b200: unknown error";
  }
}

extern "C" size_t aether_workspace_bytes(const AetherDims* dims, uint32_t flags) {
  if (!dims_ok(dims)) return 0;
  return carve(*dims, flags, nullptr).bytes + 256;
}

extern "C" size_t aether_backward_scratch_bytes(const AetherDims* dims, uint32_t flags) {
  if (!dims_ok(dims)) return 0;
  return carve_scratch(*dims, flags, nullptr).bytes + 256;
}

extern "C" int aether_caller_prep(const float* x, const float* vel, const float* charges, int32_t num_dims,
                                  const int64_t* send, const int64_t* recv, int64_t n_edges, int64_t n_nodes, float* h,
                                  float* edge_attr_orig, void* stream) {
  if (num_dims != 2 && num_dims != 3) return AETHER_E_BADARG;
  return launch_caller_prep(num_dims, x, vel, charges, send, recv, n_edges, n_nodes, h, edge_attr_orig,
                            (cudaStream_t)stream);
}

extern "C" int aether_step_forward(const AetherDims* dims, const float* const* params, const AetherGraph* graph,
                                   const float* x, const float* vel, const float* charges, const float* edge_attr_orig,
                                   const float* dropout_mask, float* out, void* workspace, size_t workspace_bytes,
                                   uint32_t flags, void* stream) {
  if (!dims_ok(dims) || !params || !graph || !out) return AETHER_E_BADARG;
  if (!workspace || ((uintptr_t)workspace & 255) || workspace_bytes < carve(*dims, flags, nullptr).bytes)
    return AETHER_E_WORKSPACE;
  const Workspace w = carve(*dims, flags, workspace);
  return forward(*dims, params, *graph, x, vel, charges, edge_attr_orig, dropout_mask, out, w, flags,
                 (cudaStream_t)stream);
}

extern "C" int aether_step_backward(const AetherDims* dims, const float* const* params, const AetherGraph* graph,
                                    const float* x, const float* vel, const float* charges,
                                    const float* edge_attr_orig, const float* dropout_mask, const float* grad_out,
                                    float* const* param_grads, float* grad_x, float* grad_vel, void* workspace,
                                    size_t workspace_bytes, void* scratch, size_t scratch_bytes, uint32_t flags,
                                    void* stream) {
  (void)edge_attr_orig;
  if (!dims_ok(dims) || !params || !graph || !grad_out || !param_grads) return AETHER_E_BADARG;
  if (!(flags & AETHER_F_SAVE_FOR_BACKWARD)) return AETHER_E_BADARG;
  if (!workspace || ((uintptr_t)workspace & 255) || workspace_bytes < carve(*dims, flags, nullptr).bytes)
    return AETHER_E_WORKSPACE;
  if (!scratch || ((uintptr_t)scratch & 255) || scratch_bytes < carve_scratch(*dims, flags, nullptr).bytes)
    return AETHER_E_WORKSPACE;
  const Workspace w = carve(*dims, flags, workspace);
  const Scratch s = carve_scratch(*dims, flags, scratch);
  return backward(*dims, params, *graph, x, vel, charges, dropout_mask, grad_out, param_grads, grad_x, grad_vel, w, s,
                  flags, (cudaStream_t)stream);
}

extern "C" int aether_part_forward_phase(const AetherDims* dims, const float* const* params, const AetherGraph* graph,
                                         const float* x, const float* vel, const float* charges,
                                         const float* edge_attr_orig, int64_t own_begin, int64_t own_end, int32_t phase,
                                         const void* xchg_in, void* xchg_out, float* out, void* workspace, size_t workspace_bytes, uint32_t flags,
                                         void* stream) {
  if (!dims_ok(dims) || !params || !graph || phase < 0 || phase > 3 || own_begin < 0 || own_end < own_begin ||
      own_end > dims->n_nodes || (phase > 0 && !xchg_in) || (phase < 3 && !xchg_out) || (phase == 3 && !out) || (flags & AETHER_F_SAVE_FOR_BACKWARD))
    return AETHER_E_BADARG;
  if (!workspace || ((uintptr_t)workspace & 255) || workspace_bytes < carve(*dims, flags, nullptr).bytes)
    return AETHER_E_WORKSPACE;
  const Workspace w = carve(*dims, flags, workspace);
  Part pt;
  pt.n0 = own_begin; pt.n1 = own_end; pt.phase_lo = pt.phase_hi = phase; pt.xchg_in = xchg_in; pt.xchg_out = xchg_out;
  return forward(*dims, params, *graph, x, vel, charges, edge_attr_orig, nullptr, out, w, flags, (cudaStream_t)stream, pt);
}

namespace aether {
namespace {
__global__ void silu_copy_kernel(const float* __restrict__ src, float* __restrict__ dst, size_t n, int apply) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = apply ? silu_f(src[i]) : src[i];
}
__global__ void strided_copy_kernel(const float* __restrict__ src, int64_t lds, float* __restrict__ dst, int64_t rows,
                                    int cols) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (size_t)rows * cols) dst[i] = src[(i / cols) * lds + i % cols];
}
__global__ void silu_copy_bf16_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, size_t n, int apply) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { const float v = __bfloat162float(src[i]); dst[i] = apply ? silu_f(v) : v; }
}
// feat16 [E, 64] = [hi(32) | lo(32)] -> fp32 [E, Fe]
__global__ void hilo_copy_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, int64_t rows, int cols) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (size_t)rows * cols) {
    const size_t r = i / cols, c = i % cols;
    dst[i] = __bfloat162float(src[r * 64 + c]) + __bfloat162float(src[r * 64 + 32 + c]);
  }
}
__global__ void sub_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ c, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) c[i] = a[i] - b[i];
}
}  // namespace
}  // namespace aether

extern "C" int aether_rollout_step(const AetherDims* dims, const float* const* params, const AetherGraph* graph,
                                   const int64_t* send64, const int64_t* recv64, const float* x, const float* vel,
                                   const float* charges, float* x_next, float* vel_next, float* edge_attr_scratch,
                                   void* workspace, size_t workspace_bytes, uint32_t flags, void* stream) {
  if (!dims_ok(dims) || !x_next || !vel_next || x_next == x || vel_next == vel) return AETHER_E_BADARG;
  if (dims->n_edges > 0 && (!send64 || !recv64 || !edge_attr_scratch)) return AETHER_E_BADARG;
  cudaStream_t st = (cudaStream_t)stream;
  AE_CHECK(launch_caller_prep(dims->num_dims, x, vel, charges, send64, recv64, dims->n_edges, dims->n_nodes, nullptr,
                              edge_attr_scratch, st));
  AE_CHECK(aether_step_forward(dims, params, graph, x, vel, charges, edge_attr_scratch, nullptr, x_next, workspace,
                               workspace_bytes, flags & ~AETHER_F_SAVE_FOR_BACKWARD, stream));
  const size_t n = (size_t)dims->n_nodes * dims->num_dims;
  if (n) sub_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x_next, x, vel_next, n);
  AE_LAUNCH_CHECK();
  return 0;
}

extern "C" int64_t aether_debug_stage(const AetherDims* dims, const void* workspace, uint32_t flags,
                                      const char* stage, float* dst, int64_t dst_floats, void* stream) {
  if (!dims_ok(dims) || !workspace || !stage || !dst) return AETHER_E_BADARG;
  const Workspace w = carve(*dims, flags, const_cast<void*>(workspace));
  const int64_t Nn = dims->n_nodes, E = dims->n_edges;
  const int D = dims->num_dims, H = dims->hidden;
  cudaStream_t st = (cudaStream_t)stream;
  const float* src = nullptr;
  int64_t n = 0;
  int silu = 0;
  if (!strcmp(stage, "field")) { src = w.f; n = Nn * D; }
  else if (!strcmp(stage, "rel_feat")) { src = w.x0; n = Nn * 3 * D; }
  else if (!strcmp(stage, "R")) { src = w.R; n = Nn * D * D; }
  else if (!strcmp(stage, "pred")) { src = w.pred; n = Nn * D; }
  else if (!strcmp(stage, "edge_attr")) {
    const int Fe = num_edge_feat(D);
    n = E * Fe;
    if (dst_floats < n) return AETHER_E_WORKSPACE;
    if (n && w.feat16) hilo_copy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const __nv_bfloat16*)w.feat16, dst, E, Fe);
    else if (n) strided_copy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(w.feat, edge_feat_ld(D), dst, E, Fe);
    return cudaPeekAtLastError() == cudaSuccess ? n : AETHER_E_BADARG;
  } else if ((stage[0] == 'e' || stage[0] == 'x') && stage[1] >= '1' && stage[1] <= '4' && !stage[2]) {
    const int l = stage[1] - '0';
    if (stage[0] == 'e') {
      src = w.pre2[l]; n = E * H; silu = 1;
      if (w.bpre2[l]) {
        if (dst_floats < n) return AETHER_E_WORKSPACE;
        if (n) silu_copy_bf16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const __nv_bfloat16*)w.bpre2[l], dst, (size_t)n, 1);
        return cudaPeekAtLastError() == cudaSuccess ? n : AETHER_E_BADARG;
      }
    }
    else { src = w.xl[l]; n = Nn * H; }
  } else return AETHER_E_BADARG;
  if (dst_floats < n) return AETHER_E_WORKSPACE;
  if (n) silu_copy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(src, dst, (size_t)n, silu);
  return cudaPeekAtLastError() == cudaSuccess ? n : AETHER_E_BADARG;
}
