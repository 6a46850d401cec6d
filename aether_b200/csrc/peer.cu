// All-gather of one exchange buffer by producer-side stores into peer memory (the receiver-partitioned
// scene's exchange of the sender term P_s / the new positions; the reference is single-device and has no collective).
//
// Every rank holds the same allocation (symmetric memory: peer_base[r] is rank r's copy, mapped into this process).
// Rank r has written block r of the buffer locally.  ONE kernel per rank then
//   1. copies its block into the same place of every peer's copy (16-byte stores over NVLink, peers taken in rotated
//      order so that the ranks do not all hit the same target at once),
//   2. after the last CTA's copies are fenced at system scope, raises flag[r] = epoch in every peer's flag row,
//   3. waits until its own flag row shows `epoch` from every peer -- their blocks have then landed here.
// No NCCL call, no host involvement; the launch is capturable (the epoch lives in device memory and advances by one per
// call, identically on every rank because all ranks issue the same sequence of exchanges).  A peer that never arrives
// makes the wait give up after ~4 s of spinning and sets state[2] (checked by the host side after a synchronise)
// instead of hanging the GPU.
#include "common.cuh"

namespace aether {
namespace {

constexpr int PEER_MAX = 16;
struct PeerPtrs { char* base[PEER_MAX]; };

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <typename V>
__device__ __forceinline__ void push_block(const PeerPtrs& pp, int world, int rank, size_t off, size_t bytes) {
  const V* src = reinterpret_cast<const V*>(pp.base[rank] + off);
  const size_t n = bytes / sizeof(V);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const V v = src[i];
    for (int d = 1; d < world; ++d) {
      const int p = (rank + d) % world;
      reinterpret_cast<V*>(pp.base[p] + off)[i] = v;
    }
  }
}

// state: [0] epoch, [1] CTAs done, [2] error (a peer did not arrive)
__global__ void __launch_bounds__(256) peer_allgather_kernel(PeerPtrs pp, int world, int rank, size_t buf_off,
                                                             size_t block_bytes, size_t flags_off, unsigned* state,
                                                             long long spin_cycles) {
  const size_t off = buf_off + (size_t)rank * block_bytes;
  if (((off | block_bytes) & 15) == 0) push_block<uint4>(pp, world, rank, off, block_bytes);
  else push_block<unsigned>(pp, world, rank, off, block_bytes);
  __threadfence_system();
  __syncthreads();
  __shared__ unsigned s_last, s_epoch;
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(&state[1], 1u);
    s_last = (prev == gridDim.x - 1);
    if (s_last) {
      __threadfence_system();
      s_epoch = state[0] + 1;
      state[0] = s_epoch;
      state[1] = 0;
    }
  }
  __syncthreads();
  if (!s_last) return;
  const unsigned epoch = s_epoch;
  const int p = threadIdx.x;
  if (p < world && p != rank) {
    st_release_sys(reinterpret_cast<unsigned*>(pp.base[p] + flags_off) + rank, epoch);
    const unsigned* mine = reinterpret_cast<const unsigned*>(pp.base[rank] + flags_off) + p;
    const long long t0 = clock64();
    // epochs only grow; the signed difference keeps the comparison right across a 32-bit wrap
    while ((int)(ld_acquire_sys(mine) - epoch) < 0) {
      if (clock64() - t0 > spin_cycles) { atomicExch(&state[2], 1u); break; }
      __nanosleep(64);
    }
  }
  __syncthreads();
  __threadfence_system();
}

}  // namespace
}  // namespace aether

using namespace aether;

extern "C" int aether_peer_allgather(void* const* peer_base, int32_t world, int32_t rank, size_t buf_offset,
                                     size_t block_bytes, size_t flags_offset, uint32_t* state, void* stream) {
  if (peer_base == nullptr || state == nullptr || world < 1 || world > PEER_MAX || rank < 0 || rank >= world)
    return AETHER_E_BADARG;
  if ((buf_offset & 3) || (block_bytes & 3) || (flags_offset & 3)) return AETHER_E_BADARG;
  if (world == 1) return 0;
  PeerPtrs pp;
  for (int i = 0; i < PEER_MAX; ++i) pp.base[i] = i < world ? static_cast<char*>(peer_base[i]) : nullptr;
  for (int i = 0; i < world; ++i)
    if (pp.base[i] == nullptr) return AETHER_E_BADARG;
  const size_t units = block_bytes / 16 + 1;
  int grid = (int)((units + 255) / 256);
  grid = grid < 1 ? 1 : (grid > 128 ? 128 : grid);        // <= one CTA per SM: every CTA is resident, the last one waits
  peer_allgather_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(pp, world, rank, buf_offset, block_bytes, flags_offset,
                                                                state, 8000000000LL);
  AE_LAUNCH_CHECK();
  return 0;
}
