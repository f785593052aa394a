// K1-K6: collectives as kernels over NVLink peer memory (symmetric arenas + signal pads).
//
// The delegates go through NCCL (ProcessGroupNCCL for DDP, torch.cuda.nccl for DataParallel) and wrap
// every collective in extra passes: copy+divide into a flat bucket, all-reduce, copy out, then a separate
// un-scale/inf-check kernel for amp (reducer.hpp:276-286, torch:amp/grad_scaler.py:349-361).  Here every
// rank's gradient arena is mapped into every other rank's address space (torch symmetric memory does the
// allocation and handle exchange; the kernels are ours) and one kernel does, per bucket:
//
//     barrier  ->  reduce-scatter by P2P loads (rank r sums shard r of all ranks)
//              ->  x 1/world, x 1/loss_scale, non-finite check          (fused epilogue)
//              ->  all-gather by P2P stores (or NVLS multimem.st)        ->  barrier
//
// in place on the gradients, so DDP needs no bucket copies at all.  With a multicast mapping the
// reduce-scatter is a single `multimem.ld_reduce` per 16 bytes: the NVSwitch adds the eight copies.
//
// Cross-GPU synchronisation: one 32-bit slot per (block, peer) in the symmetric signal pad, used as a monotonic
// meeting counter (comm/peer_sync.cuh): a meeting costs one one-way release store per peer plus a local spin.
// Every spin is bounded by the global timer and traps instead of hanging the box.
//
// Algorithm per payload size (chosen by the caller, parallel/symm.py): up to 256 KB the ONE-SHOT kernel (every rank
// reads every copy and keeps the sum: one NVLink latency, no all-gather), above it the two-shot kernel (reduce-scatter
// + all-gather, through the NVLS multicast address when there is one).
#include <cstdlib>

#include "common.cuh"
#include "comm/peer_sync.cuh"

namespace pdb {

struct PeerPtrs {
  float* buf[kMaxWorld];
  uint32_t* pad[kMaxWorld];
  float* flag[kMaxWorld];       // per-rank non-finite flag (symmetric); may be null
};

__device__ __forceinline__ float4 ld_peer(const float* p) {
  float4 v;
  // peer data is produced by another GPU: bypass L1 (.cg) so a stale line can never be observed
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_peer(float* p, const float4& v) {
  asm volatile("st.global.cg.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 multimem_ld_reduce(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// Two-shot in-place all-reduce of elements [start, start + n) of every rank's buffer.
// kMulticast: reduce and broadcast through the NVLS multicast address `mc` instead of per-peer loads/stores.
// trace (nullable): block 0 stamps the global timer at entry, after the first meeting and at exit — inside a training
// step the first interval is the wait for the slowest rank's backward, the second is the collective itself.
// kSmall: 256 threads, <= 80 registers (3 CTAs' worth of the register file), 4 requests in flight per thread.  A
// persistent convolution CTA leaves 27 K registers and ~1,700 thread slots of its SM unused; the 512-thread / 128-register
// variant needs a WHOLE register file, so behind a convolution it waits for the kernel boundary, while the small variant
// can start next to it (PDB200_COMM_SMALL_CTA=1, with twice the CTAs).
template <bool kMulticast, bool kSmall = false>
__global__ void __launch_bounds__(kSmall ? 256 : kCommThreads, kSmall ? 3 : 1) allreduce_kernel(PeerPtrs pp, float* mc, int rank, int world,
                                                                 int64_t start, int64_t n, float scale,
                                                                 const float* __restrict__ inv_scale,
                                                                 unsigned long long* __restrict__ trace) {
  if (trace && blockIdx.x == 0 && threadIdx.x == 0) trace[0] = globaltimer_ns();
  cross_rank_barrier(pp.pad, rank, world);            // everybody's gradients for this range are final
  if (trace && blockIdx.x == 0 && threadIdx.x == 0) trace[1] = globaltimer_ns();
  const float s = scale * (inv_scale ? *inv_scale : 1.0f);
  // shard r = [lo, hi) in units of float4, split evenly
  const int64_t n4 = n / 4;
  const int64_t per = (n4 + world - 1) / world;
  const int64_t lo = min(per * rank, n4), hi = min(lo + per, n4);
  bool bad = false;
  // Latency-bound over NVLink (~2 us per remote access): keep kUnroll independent 16-byte requests in
  // flight per thread (32 CTAs x 512 threads x 8 x 16 B = 2 MB, the bandwidth-delay product of one
  // NVLink direction), issue all loads of a batch before the first add.
  constexpr int kUnroll = kSmall ? 4 : 8;
  const int64_t tile = (int64_t)blockDim.x * kUnroll;
  for (int64_t base = lo + (int64_t)blockIdx.x * tile; base < hi; base += (int64_t)gridDim.x * tile) {
    float4 acc[kUnroll];
    int64_t off[kUnroll];
    bool live[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
      live[u] = i < hi;
      off[u] = start + (live[u] ? i : lo) * 4;
    }
    if (kMulticast) {
#pragma unroll
      for (int u = 0; u < kUnroll; ++u)
        if (live[u]) acc[u] = multimem_ld_reduce(mc + off[u]);
    } else {
#pragma unroll
      for (int u = 0; u < kUnroll; ++u)
        if (live[u]) acc[u] = ld_peer(pp.buf[rank] + off[u]);
#pragma unroll 1
      for (int k = 1; k < world; ++k) {
        const int p = rank + k < world ? rank + k : rank + k - world;     // stagger peers across ranks
        const float* src = pp.buf[p];
        float4 v[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
          if (live[u]) v[u] = ld_peer(src + off[u]);
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
          if (live[u]) acc[u].x += v[u].x, acc[u].y += v[u].y, acc[u].z += v[u].z, acc[u].w += v[u].w;
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (!live[u]) continue;
      acc[u].x *= s, acc[u].y *= s, acc[u].z *= s, acc[u].w *= s;
      bad |= !(isfinite(acc[u].x) && isfinite(acc[u].y) && isfinite(acc[u].z) && isfinite(acc[u].w));
    }
    if (kMulticast) {
#pragma unroll
      for (int u = 0; u < kUnroll; ++u)
        if (live[u]) multimem_st(mc + off[u], acc[u]);
    } else {
#pragma unroll 1
      for (int k = 0; k < world; ++k) {
        const int p = rank + k < world ? rank + k : rank + k - world;
        float* dst = pp.buf[p];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
          if (live[u]) st_peer(dst + off[u], acc[u]);
      }
    }
  }
  if (pp.flag[0] != nullptr && __syncthreads_or(bad) && threadIdx.x < world) {
    // the non-finite verdict must be identical on all ranks: raise it everywhere
    *reinterpret_cast<volatile float*>(pp.flag[threadIdx.x]) = 1.0f;
  }
  __threadfence_system();
  cross_rank_barrier(pp.pad, rank, world);            // all shards have landed everywhere
  if (trace && blockIdx.x == 0 && threadIdx.x == 0) trace[2] = globaltimer_ns();
}

// ---- 16-bit wire variant (fp16_compress / bf16_compress hooks, default_comm_hooks.hpp:9-34) -------------------
// Gradients stay fp32 in the arena; only what crosses NVLink is 16-bit:
//   pack own range -> S (x 1/world, like the hook's div)  | barrier |  rank r: sum shard r of every S in fp32,
//   x 1/loss_scale, non-finite check, store 16-bit into every R  | barrier |  unpack R -> fp32 gradients.
// S and R are the two halves of a symmetric staging buffer (elements of W, `half_elems` each).
struct PeerPtrs16 {
  void* stage[kMaxWorld];
};

template <typename W>
__global__ void __launch_bounds__(kCommThreads) allreduce_wire16_kernel(PeerPtrs pp, PeerPtrs16 st, int rank, int world,
                                                                        int64_t start, int64_t n, int64_t half_elems,
                                                                        float pre_scale,
                                                                        const float* __restrict__ inv_scale) {
  const int64_t n8 = n / 8;
  float* grads = pp.buf[rank] + start;
  W* S_own = static_cast<W*>(st.stage[rank]);
  W* R_own = S_own + half_elems;
  constexpr int kUnroll = 4;
  const int64_t tile = (int64_t)blockDim.x * kUnroll;
  const int64_t per = (n8 + world - 1) / world;
  // The cross-rank barrier pairs block b of this rank with block b of every peer — nothing orders block b
  // against the OTHER blocks of a peer.  So block b packs (phase 1) and unpacks (phase 3) exactly the vectors
  // that block b of the shard owner reads / writes in phase 2: the same index walk, repeated for every shard.
  for (int q = 0; q < world; ++q) {                    // phase 1: pack (local)
    const int64_t qlo = min(per * q, n8), qhi = min(qlo + per, n8);
    for (int64_t base = qlo + (int64_t)blockIdx.x * tile; base < qhi; base += (int64_t)gridDim.x * tile) {
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
        if (i >= qhi) continue;
        const float4 a = reinterpret_cast<const float4*>(grads)[2 * i], b = reinterpret_cast<const float4*>(grads)[2 * i + 1];
        float f[8] = {a.x * pre_scale, a.y * pre_scale, a.z * pre_scale, a.w * pre_scale,
                      b.x * pre_scale, b.y * pre_scale, b.z * pre_scale, b.w * pre_scale};
        reinterpret_cast<uint4*>(S_own)[i] = pack8<W>(f);
      }
    }
  }
  __threadfence_system();
  cross_rank_barrier(pp.pad, rank, world);
  const float s = inv_scale ? *inv_scale : 1.0f;
  const int64_t lo = min(per * rank, n8), hi = min(lo + per, n8);
  bool bad = false;
  for (int64_t base = lo + (int64_t)blockIdx.x * tile; base < hi; base += (int64_t)gridDim.x * tile) {
    float acc[kUnroll][8];
    bool live[kUnroll];
    int64_t idx[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      idx[u] = base + (int64_t)u * blockDim.x + threadIdx.x;
      live[u] = idx[u] < hi;
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[u][k] = 0.f;
    }
#pragma unroll 1
    for (int kk = 0; kk < world; ++kk) {
      const int p = rank + kk < world ? rank + kk : rank + kk - world;
      const uint4* src = static_cast<const uint4*>(st.stage[p]);
      uint4 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u)
        if (live[u])
          asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + idx[u]));
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        if (!live[u]) continue;
        float f[8];
        unpack8<W>(v[u], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[u][k] += f[k];
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (!live[u]) continue;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        acc[u][k] *= s;
        bad |= !isfinite(acc[u][k]);
      }
      const uint4 w = pack8<W>(acc[u]);
#pragma unroll 1
      for (int kk = 0; kk < world; ++kk) {
        const int p = rank + kk < world ? rank + kk : rank + kk - world;
        uint4* dst = reinterpret_cast<uint4*>(static_cast<W*>(st.stage[p]) + half_elems) + idx[u];
        asm volatile("st.global.cg.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "r"(w.x), "r"(w.y), "r"(w.z), "r"(w.w) : "memory");
      }
    }
  }
  if (pp.flag[0] != nullptr && __syncthreads_or(bad) && threadIdx.x < world)
    *reinterpret_cast<volatile float*>(pp.flag[threadIdx.x]) = 1.0f;
  __threadfence_system();
  cross_rank_barrier(pp.pad, rank, world);
  for (int q = 0; q < world; ++q) {                    // phase 3: unpack (local; shard q was written by rank q's block b)
    const int64_t qlo = min(per * q, n8), qhi = min(qlo + per, n8);
    for (int64_t base = qlo + (int64_t)blockIdx.x * tile; base < qhi; base += (int64_t)gridDim.x * tile) {
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
        if (i >= qhi) continue;
        uint4 v;
        asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(reinterpret_cast<const uint4*>(R_own) + i));
        float f[8];
        unpack8<W>(v, f);
        reinterpret_cast<float4*>(grads)[2 * i] = make_float4(f[0], f[1], f[2], f[3]);
        reinterpret_cast<float4*>(grads)[2 * i + 1] = make_float4(f[4], f[5], f[6], f[7]);
      }
    }
  }
}

// Broadcast of [start, start+n) from `root` to every rank: non-root ranks pull over NVLink.
__global__ void __launch_bounds__(kCommThreads) broadcast_kernel(PeerPtrs pp, int rank, int world, int root,
                                                                 int64_t start, int64_t n) {
  cross_rank_barrier(pp.pad, rank, world);
  if (rank != root) {
    const int64_t n4 = n / 4, stride = (int64_t)gridDim.x * blockDim.x;
    const float* src = pp.buf[root] + start;
    float* dst = pp.buf[rank] + start;
    constexpr int kUnroll = 8;
    const int64_t tile = (int64_t)blockDim.x * kUnroll;
    (void)stride;
    for (int64_t base = (int64_t)blockIdx.x * tile; base < n4; base += (int64_t)gridDim.x * tile) {
      float4 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
        if (i < n4) v[u] = ld_peer(src + i * 4);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
        if (i < n4) reinterpret_cast<float4*>(dst)[i] = v[u];
      }
    }
  }
  cross_rank_barrier(pp.pad, rank, world);            // root may overwrite its copy only after everyone has read it
}

// One-shot IN-PLACE all-reduce of a small range (gradient buckets up to 256 KB: shrunken models, generic modules):
// meet, every rank reads all `world` copies and adds them in rank order (so every rank holds bit-identical sums and
// the non-finite verdict is the same everywhere without a flag exchange), meet again (nobody may overwrite a copy
// that a peer is still reading), write the local copy.  The whole range lives in registers between the two
// meetings: gridDim.x * blockDim.x * kOneShotUnroll float4 must cover it (the launcher sizes the grid).
constexpr int kOneShotUnroll = 4;
__global__ void __launch_bounds__(kCommThreads) allreduce_oneshot_kernel(PeerPtrs pp, int rank, int world, int64_t start,
                                                                         int64_t n, float scale,
                                                                         const float* __restrict__ inv_scale) {
  cross_rank_barrier(pp.pad, rank, world);
  const float s = scale * (inv_scale ? *inv_scale : 1.0f);
  const int64_t n4 = n / 4;
  float4 acc[kOneShotUnroll];
  int64_t idx[kOneShotUnroll];
  bool bad = false;
#pragma unroll
  for (int u = 0; u < kOneShotUnroll; ++u) {
    idx[u] = ((int64_t)blockIdx.x * kOneShotUnroll + u) * blockDim.x + threadIdx.x;
    acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll 1
  for (int p = 0; p < world; ++p) {
    const float* src = pp.buf[p] + start;
    float4 v[kOneShotUnroll];
#pragma unroll
    for (int u = 0; u < kOneShotUnroll; ++u)
      if (idx[u] < n4) v[u] = ld_peer(src + idx[u] * 4);
#pragma unroll
    for (int u = 0; u < kOneShotUnroll; ++u)
      if (idx[u] < n4) acc[u].x += v[u].x, acc[u].y += v[u].y, acc[u].z += v[u].z, acc[u].w += v[u].w;
  }
#pragma unroll
  for (int u = 0; u < kOneShotUnroll; ++u) {
    if (idx[u] >= n4) continue;
    acc[u].x *= s, acc[u].y *= s, acc[u].z *= s, acc[u].w *= s;
    bad |= !(isfinite(acc[u].x) && isfinite(acc[u].y) && isfinite(acc[u].z) && isfinite(acc[u].w));
  }
  cross_rank_barrier(pp.pad, rank, world);            // every rank has read every copy
  float* dst = pp.buf[rank] + start;
#pragma unroll
  for (int u = 0; u < kOneShotUnroll; ++u)
    if (idx[u] < n4) reinterpret_cast<float4*>(dst)[idx[u]] = acc[u];
  if (pp.flag[0] != nullptr && __syncthreads_or(bad) && threadIdx.x == 0)
    *reinterpret_cast<volatile float*>(pp.flag[rank]) = 1.0f;      // same sums on every rank: a local flag suffices
}

// All-reduce of a tiny fp32 vector that is NOT in symmetric memory (SyncBN statistics outside the fused kernels,
// metric vectors): one block per rank, one meeting (exchange_sum, comm/peer_sync.cuh).
// staging layout per rank: float[2][world][slot_stride].
__global__ void __launch_bounds__(1024) small_allreduce_kernel(SyncPeers sp, float* __restrict__ data, int n) {
  exchange_sum(sp, data, data, n, threadIdx.x, blockDim.x, [] { __syncthreads(); });
}

// ---- single-process DataParallel helpers (all GPUs in one address space, peer access enabled) ----------
__global__ void __launch_bounds__(256) peer_pull_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src, int64_t n16) {
  constexpr int kUnroll = 8;
  const int64_t tile = (int64_t)blockDim.x * kUnroll;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < n16; base += (int64_t)gridDim.x * tile) {
    uint4 v[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
      if (i < n16)
        asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + i));
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
      if (i < n16) dst[i] = v[u];
    }
  }
}

struct PeerList {
  const float* p[kMaxWorld];
};

// Shard-wise gather (single process, all GPUs in one address space): shard q of the destination array is copied from
// the same byte range of source q.  Second half of the sharded parameter replicate / gradient reduce of the
// DataParallel mode: instead of GPU 0 sending (or receiving) n-1 full arenas, every GPU first moves / reduces ONE shard
// and then all shards are exchanged peer to peer, so GPU 0's NVLink carries ~2/n of the bytes it carried before.
__global__ void __launch_bounds__(256) peer_gather_kernel(uint4* __restrict__ dst, PeerList srcs, int64_t shard16,
                                                          int64_t total16, int skip) {
  const int q = blockIdx.y;
  if (q == skip || srcs.p[q] == nullptr) return;
  const int64_t lo = (int64_t)q * shard16, hi = min(lo + shard16, total16);
  const uint4* src = reinterpret_cast<const uint4*>(srcs.p[q]);
  constexpr int kUnroll = 8;
  const int64_t tile = (int64_t)blockDim.x * kUnroll;
  for (int64_t base = lo + (int64_t)blockIdx.x * tile; base < hi; base += (int64_t)gridDim.x * tile) {
    uint4 v[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
      if (i < hi)
        asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + i));
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
      if (i < hi) dst[i] = v[u];
    }
  }
}

__global__ void __launch_bounds__(256) reduce_to_root_kernel(float* __restrict__ root, PeerList peers, int n_peers,
                                                             int64_t n4) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 acc = reinterpret_cast<float4*>(root)[i];
    for (int k = 0; k < n_peers; ++k) {
      const float4 v = ld_peer(peers.p[k] + i * 4);
      acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
    }
    reinterpret_cast<float4*>(root)[i] = acc;
  }
}

static int fill_peers(PeerPtrs* pp, void* const* bufs, void* const* pads, void* const* flags, int world) {
  PDB_CHECK_ARG(world >= 1 && world <= kMaxWorld, "world size %d not supported (max %d)", world, kMaxWorld);
  for (int i = 0; i < kMaxWorld; ++i) pp->buf[i] = nullptr, pp->pad[i] = nullptr, pp->flag[i] = nullptr;
  for (int i = 0; i < world; ++i) {
    pp->buf[i] = static_cast<float*>(bufs[i]);
    pp->pad[i] = static_cast<uint32_t*>(pads[i]);
    pp->flag[i] = flags ? static_cast<float*>(flags[i]) : nullptr;
  }
  return 0;
}

}  // namespace pdb
using namespace pdb;

// --profile: a device buffer of 3 x capacity uint64 that successive pdb_allreduce_inplace launches stamp (see
// allreduce_kernel); null switches tracing off.  Returns the number of launches traced so far into the previous buffer.
static unsigned long long* g_trace = nullptr;
static int g_trace_cap = 0, g_trace_next = 0;
extern "C" int pdb_comm_set_trace(void* buf, int capacity) {
  const int used = g_trace_next;
  g_trace = static_cast<unsigned long long*>(buf);
  g_trace_cap = buf ? capacity : 0;
  g_trace_next = 0;
  return used;
}

extern "C" int pdb_comm_max_blocks(int world, size_t signal_pad_bytes) {
  const size_t slots = signal_pad_bytes / sizeof(uint32_t);
  int b = (int)(slots / (size_t)(world > 0 ? world : 1));
  return b > 64 ? 64 : b;
}

extern "C" int pdb_allreduce_inplace(void* const* bufs, void* const* pads, void* const* flags, void* multicast, int rank,
                                     int world, int64_t start, int64_t n, float scale, const float* inv_scale,
                                     int blocks, void* stream) {
  PDB_CHECK_ARG(bufs && pads && n >= 0 && start >= 0, "pdb_allreduce_inplace: bad arguments");
  PDB_CHECK_ARG(n % 4 == 0 && start % 4 == 0, "pdb_allreduce_inplace: range must be 16-byte aligned (start=%lld n=%lld)",
                (long long)start, (long long)n);
  PDB_CHECK_ARG(rank >= 0 && rank < world, "pdb_allreduce_inplace: rank %d out of range", rank);
  PDB_CHECK_ARG(blocks >= 1 && blocks <= 64, "pdb_allreduce_inplace: blocks must be in [1, 64]");
  if (n == 0) return 0;
  PeerPtrs pp;
  if (int e = fill_peers(&pp, bufs, pads, flags, world)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long* tr = nullptr;
  if (g_trace && g_trace_next < g_trace_cap) tr = g_trace + 3 * (g_trace_next++);
  static const bool small_cta = [] { const char* e = getenv("PDB200_COMM_SMALL_CTA"); return e && e[0] == '1'; }();
  if (small_cta) {
    if (multicast)
      allreduce_kernel<true, true><<<blocks, 256, 0, st>>>(pp, static_cast<float*>(multicast), rank, world, start, n, scale, inv_scale, tr);
    else
      allreduce_kernel<false, true><<<blocks, 256, 0, st>>>(pp, nullptr, rank, world, start, n, scale, inv_scale, tr);
  } else if (multicast)
    allreduce_kernel<true><<<blocks, kCommThreads, 0, st>>>(pp, static_cast<float*>(multicast), rank, world, start, n, scale, inv_scale, tr);
  else
    allreduce_kernel<false><<<blocks, kCommThreads, 0, st>>>(pp, nullptr, rank, world, start, n, scale, inv_scale, tr);
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_allreduce_inplace_wire16(void* const* bufs, void* const* pads, void* const* flags, void* const* staging,
                                            int64_t staging_half_elems, int wire_dtype, int rank, int world,
                                            int64_t start, int64_t n, float pre_scale, const float* inv_scale,
                                            int blocks, void* stream) {
  PDB_CHECK_ARG(bufs && pads && staging && n >= 0 && start >= 0, "pdb_allreduce_inplace_wire16: bad arguments");
  PDB_CHECK_ARG(n % 8 == 0 && start % 4 == 0, "pdb_allreduce_inplace_wire16: n must be a multiple of 8, start of 4");
  PDB_CHECK_ARG(n <= staging_half_elems, "pdb_allreduce_inplace_wire16: range (%lld) exceeds the staging half (%lld)",
                (long long)n, (long long)staging_half_elems);
  PDB_CHECK_ARG(rank >= 0 && rank < world && blocks >= 1 && blocks <= 64, "pdb_allreduce_inplace_wire16: bad rank/blocks");
  if (n == 0) return 0;
  PeerPtrs pp;
  if (int e = fill_peers(&pp, bufs, pads, flags, world)) return e;
  PeerPtrs16 st16;
  for (int i = 0; i < kMaxWorld; ++i) st16.stage[i] = i < world ? staging[i] : nullptr;
  cudaStream_t st = (cudaStream_t)stream;
  PDB_DISPATCH_16(wire_dtype, W, (allreduce_wire16_kernel<W><<<blocks, kCommThreads, 0, st>>>(
                                     pp, st16, rank, world, start, n, staging_half_elems, pre_scale, inv_scale)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_broadcast_inplace(void* const* bufs, void* const* pads, int rank, int world, int root, int64_t start,
                                     int64_t n, int blocks, void* stream) {
  PDB_CHECK_ARG(bufs && pads && n >= 0 && n % 4 == 0 && start % 4 == 0, "pdb_broadcast_inplace: bad arguments");
  PDB_CHECK_ARG(rank >= 0 && rank < world && root >= 0 && root < world, "pdb_broadcast_inplace: bad rank/root");
  PDB_CHECK_ARG(blocks >= 1 && blocks <= 64, "pdb_broadcast_inplace: blocks must be in [1, 64]");
  if (n == 0) return 0;
  PeerPtrs pp;
  if (int e = fill_peers(&pp, bufs, pads, nullptr, world)) return e;
  broadcast_kernel<<<blocks, kCommThreads, 0, (cudaStream_t)stream>>>(pp, rank, world, root, start, n);
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_allreduce_oneshot(void* const* bufs, void* const* pads, void* const* flags, int rank, int world,
                                     int64_t start, int64_t n, float scale, const float* inv_scale, void* stream) {
  PDB_CHECK_ARG(bufs && pads && n >= 0 && start >= 0, "pdb_allreduce_oneshot: bad arguments");
  PDB_CHECK_ARG(n % 4 == 0 && start % 4 == 0, "pdb_allreduce_oneshot: range must be 16-byte aligned (start=%lld n=%lld)",
                (long long)start, (long long)n);
  PDB_CHECK_ARG(rank >= 0 && rank < world, "pdb_allreduce_oneshot: rank %d out of range", rank);
  if (n == 0) return 0;
  const int64_t per_block = (int64_t)kCommThreads * kOneShotUnroll;            // float4 per block
  const int64_t blocks = (n / 4 + per_block - 1) / per_block;
  PDB_CHECK_ARG(blocks <= 64, "pdb_allreduce_oneshot: range of %lld floats is too large for the one-shot kernel (max %lld)",
                (long long)n, (long long)(64 * per_block * 4));
  PeerPtrs pp;
  if (int e = fill_peers(&pp, bufs, pads, flags, world)) return e;
  allreduce_oneshot_kernel<<<(int)blocks, kCommThreads, 0, (cudaStream_t)stream>>>(pp, rank, world, start, n, scale, inv_scale);
  PDB_LAUNCH_CHECK();
  return 0;
}

namespace pdb {
int fill_sync(SyncPeers* sp, void* const* staging, void* const* pads, int rank, int world, int slot_stride, int row) {
  PDB_CHECK_ARG(staging && pads, "sync peers: null staging / pads");
  PDB_CHECK_ARG(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, "sync peers: bad rank %d / world %d", rank, world);
  for (int i = 0; i < kMaxWorld; ++i) {
    sp->staging[i] = i < world ? static_cast<float*>(staging[i]) : nullptr;
    sp->pad[i] = i < world ? static_cast<uint32_t*>(pads[i]) : nullptr;
  }
  sp->rank = rank, sp->world = world, sp->slot_stride = slot_stride, sp->row = row;
  return 0;
}
}  // namespace pdb

extern "C" int pdb_small_allreduce(void* const* staging, void* const* pads, int rank, int world, float* data, int n,
                                   int slot_stride, void* stream) {
  PDB_CHECK_ARG(staging && pads && data && n > 0 && n <= slot_stride, "pdb_small_allreduce: bad arguments");
  SyncPeers sp;
  if (int e = fill_sync(&sp, staging, pads, rank, world, slot_stride, 0)) return e;
  small_allreduce_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(sp, data, n);
  PDB_LAUNCH_CHECK();
  return 0;
}

static int enable_peer(int dst_dev, int src_dev) {
  if (dst_dev == src_dev) return 0;
  int can = 0;
  PDB_CUDA(cudaDeviceCanAccessPeer(&can, dst_dev, src_dev));
  PDB_CHECK_ARG(can, "GPU %d cannot access GPU %d over P2P", dst_dev, src_dev);
  cudaError_t e = cudaDeviceEnablePeerAccess(src_dev, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  PDB_CUDA(e);
  return 0;
}

static int device_of(const void* p, int* dev) {
  cudaPointerAttributes a;
  PDB_CUDA(cudaPointerGetAttributes(&a, p));
  *dev = a.device;
  return 0;
}

extern "C" int pdb_peer_pull(void* dst, const void* src, int64_t bytes, void* stream) {
  PDB_CHECK_ARG(dst && src && bytes > 0 && bytes % 16 == 0, "pdb_peer_pull: bytes (%lld) must be a positive multiple of 16",
                (long long)bytes);
  PDB_CHECK_ARG((((uintptr_t)dst | (uintptr_t)src) & 15) == 0, "pdb_peer_pull: pointers must be 16-byte aligned");
  int cur = 0, sdev = 0;
  PDB_CUDA(cudaGetDevice(&cur));
  if (int e = device_of(src, &sdev)) return e;
  if (int e = enable_peer(cur, sdev)) return e;
  const int64_t n16 = bytes / 16;
  int64_t want = (n16 + 255) / 256;
  const int grid = (int)(want < 128 ? want : 128);        // ~2 MB in flight saturates an NVLink direction
  peer_pull_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((uint4*)dst, (const uint4*)src, n16);
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_reduce_to_root(float* root, void* const* peers, int n_peers, int64_t n, void* stream) {
  PDB_CHECK_ARG(root && peers && n_peers >= 1 && n_peers < kMaxWorld && n > 0 && n % 4 == 0,
                "pdb_reduce_to_root: bad arguments");
  int cur = 0;
  PDB_CUDA(cudaGetDevice(&cur));
  PeerList pl;
  for (int i = 0; i < n_peers; ++i) {
    int d = 0;
    if (int e = device_of(peers[i], &d)) return e;
    if (int e = enable_peer(cur, d)) return e;
    pl.p[i] = static_cast<const float*>(peers[i]);
  }
  const int64_t n4 = n / 4;
  int64_t want = (n4 + 255) / 256;
  const int grid = (int)(want < 128 ? want : 128);
  reduce_to_root_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(root, pl, n_peers, n4);
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_peer_gather(void* dst, void* const* srcs, int n, int64_t shard_bytes, int64_t total_bytes, int skip,
                               void* stream) {
  PDB_CHECK_ARG(dst && srcs && n >= 1 && n <= kMaxWorld, "pdb_peer_gather: bad arguments");
  PDB_CHECK_ARG(shard_bytes > 0 && shard_bytes % 16 == 0 && total_bytes > 0 && total_bytes % 16 == 0,
                "pdb_peer_gather: shard (%lld) and total (%lld) bytes must be positive multiples of 16",
                (long long)shard_bytes, (long long)total_bytes);
  PDB_CHECK_ARG(((uintptr_t)dst & 15) == 0, "pdb_peer_gather: dst must be 16-byte aligned");
  int cur = 0;
  PDB_CUDA(cudaGetDevice(&cur));
  PeerList pl;
  for (int i = 0; i < kMaxWorld; ++i) pl.p[i] = nullptr;
  for (int i = 0; i < n; ++i) {
    if (i == skip || srcs[i] == nullptr) continue;
    PDB_CHECK_ARG(((uintptr_t)srcs[i] & 15) == 0, "pdb_peer_gather: sources must be 16-byte aligned");
    int d = 0;
    if (int e = device_of(srcs[i], &d)) return e;
    if (int e = enable_peer(cur, d)) return e;
    pl.p[i] = static_cast<const float*>(srcs[i]);
  }
  const int64_t shard16 = shard_bytes / 16;
  int64_t want = (shard16 + 2047) / 2048;
  dim3 grid((unsigned)(want < 32 ? (want > 0 ? want : 1) : 32), (unsigned)n);     // n x 32 blocks x 32 KB in flight
  peer_gather_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((uint4*)dst, pl, shard16, total_bytes / 16, skip);
  PDB_LAUNCH_CHECK();
  return 0;
}
