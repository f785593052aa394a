// Cross-GPU synchronisation primitives shared by the collective kernels, the SyncBatchNorm exchanges inside the
// convolution / BatchNorm kernels, and the small all-reduce.
//
// Signal pads: every rank owns an array of 32-bit slots that all peers can write (symmetric memory).  Row `row` of a
// pad holds one slot per source rank: pad[row * world + src].  A meeting of row `row` is
//
//     epoch = ++(my own diagonal slot pad_me[row * world + me])        (a private count of meetings held so far)
//     for every peer p:  st.release.sys  pad_p[row * world + me] = epoch      (one-way store, nobody waits for it)
//     for every peer p:  spin on ld.acquire.sys pad_me[row * world + p] until it reaches epoch
//
// Counters only grow, so nothing is reset and nothing can be consumed twice: a peer cannot run more than one meeting
// ahead (its next signal needs my next signal), so "slot >= epoch" is exact.  The first version used a compare-and-swap
// ping-pong (set the remote flag when it is 0, then swap the local one back to 0): two dependent NVLink round trips per
// meeting where this needs one one-way store.  The sequence number lives in the pad itself, not on the host, so
// kernels that are replayed by a CUDA graph keep counting.
// Every spin is bounded by the global timer and traps instead of hanging the box.
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

__device__ __forceinline__ void st_release_sys(uint32_t* addr, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* addr) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* addr) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}

// Number of the meeting of `row` that this rank holds next.  Every thread that will call meet_signal_wait reads it
// BEFORE a group-wide barrier; meet_signal_wait bumps the private counter after that barrier.
__device__ __forceinline__ uint32_t meet_next_epoch(uint32_t* const* pads, int rank, int world, int row) {
  return ld_volatile_u32(pads[rank] + row * world + rank) + 1u;
}

// The cross-rank part of a meeting: thread `tid` (0 <= tid < world) of the calling group handles peer `tid`.  The
// caller brackets this with its own group barrier (__syncthreads or a named barrier) and has made its data visible
// (__threadfence_system) before.
__device__ __forceinline__ void meet_signal_wait(uint32_t* const* pads, int rank, int world, int row, int tid,
                                                 uint32_t epoch) {
  if (tid >= world) return;
  uint32_t* mine = pads[rank] + row * world;
  if (tid == rank) {
    *reinterpret_cast<volatile uint32_t*>(mine + rank) = epoch;
    return;
  }
  st_release_sys(pads[tid] + row * world + rank, epoch);
  const unsigned long long t0 = globaltimer_ns();
  while ((int32_t)(ld_acquire_sys(mine + tid) - epoch) < 0)
    if (globaltimer_ns() - t0 > kSpinTimeoutNs) __trap();
}

// Block-level barrier across ranks: block b of every rank meets block b of every other rank (row = blockIdx.x).
__device__ __forceinline__ void cross_rank_barrier(uint32_t* const* pads, int rank, int world) {
  const uint32_t epoch = meet_next_epoch(pads, rank, world, blockIdx.x);
  __syncthreads();
  meet_signal_wait(pads, rank, world, blockIdx.x, threadIdx.x, epoch);
  __syncthreads();
}

// ---- SyncBatchNorm / small-vector exchange ---------------------------------------------------------------------------
// One block (or one synchronising group of threads) per rank holds n <= slot_stride floats that have to be summed over
// the ranks.  Staging per rank: float[2][world][slot_stride]; exchange number e uses half e & 1, so a rank may start
// writing the next exchange while a slower peer still reads this one, and ONE meeting per exchange is enough (a rank
// can only be two exchanges ahead after the peer has signalled the one in between, i.e. after it finished reading).
struct SyncPeers {
  float* staging[kMaxWorld];
  uint32_t* pad[kMaxWorld];
  int rank, world;
  int slot_stride;      // floats per rank slot
  int row;              // pad row used for the meetings
};

// vals[0:n] (global memory of this rank, written by this group before a group barrier) -> out[0:n] = sum over ranks,
// added in rank order on every rank (bit-identical results everywhere).  tid / nthreads: index and size of the calling
// group; sync() is its barrier.  vals may alias out.
template <typename Sync>
__device__ __forceinline__ void exchange_sum(const SyncPeers& sp, const float* vals, float* out, int n, int tid,
                                             int nthreads, Sync sync) {
  const uint32_t epoch = meet_next_epoch(sp.pad, sp.rank, sp.world, sp.row);
  const int64_t half = (int64_t)(epoch & 1u) * sp.world * sp.slot_stride;
  for (int c = tid; c < n; c += nthreads) {
    const float v = __ldcg(vals + c);
    for (int p = 0; p < sp.world; ++p) sp.staging[p][half + (int64_t)sp.rank * sp.slot_stride + c] = v;
  }
  __threadfence_system();
  sync();
  meet_signal_wait(sp.pad, sp.rank, sp.world, sp.row, tid, epoch);
  sync();
  for (int c = tid; c < n; c += nthreads) {
    float acc = 0.f;
    for (int p = 0; p < sp.world; ++p) acc += __ldcg(sp.staging[sp.rank] + half + (int64_t)p * sp.slot_stride + c);
    out[c] = acc;
  }
}

}  // namespace pdb
