// Cross-GPU synchronisation primitives shared by the collective kernels and the SyncBatchNorm statistics kernel.
#pragma once
#include "common.cuh"

namespace pdb {

constexpr int kMaxWorld = 16;
constexpr int kCommThreads = 512;
constexpr unsigned long long kSpinTimeoutNs = 20ull * 1000 * 1000 * 1000;     // 20 s

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.release.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}

// Block-level barrier across ranks: block b of every rank meets block b of every other rank.
// Slot layout in each rank's pad: pad[b * world + src_rank].
__device__ __forceinline__ void cross_rank_barrier(uint32_t* const* pads, int rank, int world) {
  __syncthreads();
  if (threadIdx.x < world && (int)threadIdx.x != rank) {
    const int peer = threadIdx.x;
    uint32_t* remote = pads[peer] + blockIdx.x * world + rank;        // my flag in the peer's pad
    uint32_t* local = pads[rank] + blockIdx.x * world + peer;         // the peer's flag in my pad
    const unsigned long long t0 = globaltimer_ns();
    while (cas_release_sys(remote, 0u, 1u) != 0u)
      if (globaltimer_ns() - t0 > kSpinTimeoutNs) __trap();
    while (cas_acquire_sys(local, 1u, 0u) != 1u)
      if (globaltimer_ns() - t0 > kSpinTimeoutNs) __trap();
  }
  __syncthreads();
}

}  // namespace pdb
