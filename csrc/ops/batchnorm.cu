// K8: BatchNorm (+ReLU, +residual) forward / backward on NHWC 16-bit activations, fp32 statistics.
//
// The delegate (cuDNN spatial-persistent BN + separate ATen ReLU / add kernels, torch:nn/modules/
// batchnorm.py:144-210) makes 3-4 passes over the activation per BN layer.  Here:
//   forward : statistics come out of the producing convolution's epilogue (or pdb_bn_stats), then ONE
//             pass normalises, adds the residual and applies ReLU; the per-channel finalisation
//             (mean/rstd, running-stat update) is done inside that pass's prologue, not in a launch
//             of its own;
//   backward: one reduction pass (ReLU mask folded in) and one apply pass that also emits the residual
//             branch gradient and the dgamma/dbeta accumulation.
// Layout [rows, C] with C contiguous: a thread owns 8 consecutive channels (one 128-bit access).
#include "common.cuh"
#include "bn_fast.cuh"
#include "comm/peer_sync.cuh"

namespace pdb {

int fill_sync(SyncPeers* sp, void* const* staging, void* const* pads, int rank, int world, int slot_stride, int row);   // collectives.cu

#ifndef PDB_BN_REDUCE_MINBLOCKS
#define PDB_BN_REDUCE_MINBLOCKS 3      // resident blocks per SM the backward reduction is compiled for (85 registers)
#endif
constexpr int kThreads = 256;
constexpr int kMaxC = 4096;

// ------------------------------------------------------------------------------------------------
// Channel reductions.  Block = cx channel-groups (8 ch each) x ry rows.
// MODE 0: (sum x, sum x^2)           — forward statistics
// MODE 1: (sum g, sum g*xhat), g = dz * (z>0)   — backward reductions
template <typename T, int MODE, int kMask /* 0 none, 1 z > 0, 2 bit mask */>
__device__ __forceinline__ void reduce_rows(const T* __restrict__ a, const T* __restrict__ z,
                                            const uint8_t* __restrict__ bits, const T* __restrict__ x, int64_t rows,
                                            int C, int cg, int cx, int ry,
                                            int ty, const float* __restrict__ mean, const float* __restrict__ rstd,
                                            float (&s0)[8], float (&s1)[8]) {
  // MODE 1 accumulates the RAW moments sum(g) and sum(g*x); sum(g*xhat) = rstd * (sum(g*x) - mean*sum(g)) is
  // formed once per channel when the block total is published — 16 fewer live registers per thread, which is
  // what lets three blocks share an SM (this kernel needs bytes in flight, not arithmetic)
  (void)mean;
  (void)rstd;
  constexpr int U = sizeof(T) == 4 ? 2 : 4;   // rows in flight per thread (same bytes in flight for fp32)
  // a block walks slabs of U*ry ADJACENT rows (U*ry*C*2 bytes contiguous per stream, >= 16 KB): DRAM likes
  // long contiguous bursts far better than U scattered 4 KB pieces
  const int64_t rstride = ry;
  for (int64_t r0 = (int64_t)blockIdx.x * ry * U + ty; r0 < rows; r0 += (int64_t)gridDim.x * ry * U) {
    V8<T> va[U], vx[U], vz[U];
    uint32_t vb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t r = r0 + u * rstride;
      if (r < rows) {
        const int64_t off = r * C + (int64_t)cg * 8;
        va[u] = MODE == 1 ? ld8s_split<T>(a + off) : ld8s<T>(a + off);
        if (MODE == 1) {
          vx[u] = ld8s_split<T>(x + off);
          if (kMask == 1) vz[u] = ld8s_split<T>(z + off);
          if (kMask == 2) vb[u] = __ldg(bits + (off >> 3));
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t r = r0 + u * rstride;
      if (r >= rows) continue;
      float fa[8];
      unpack8<T>(va[u], fa);
      if (MODE == 0) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          s0[k] += fa[k];
          s1[k] = fmaf(fa[k], fa[k], s1[k]);
        }
      } else {
        float fx[8];
        unpack8<T>(vx[u], fx);
        if (kMask == 1) {
          float fz[8];
          unpack8<T>(vz[u], fz);
#pragma unroll
          for (int k = 0; k < 8; ++k) fa[k] = fz[k] > 0.f ? fa[k] : 0.f;
        }
        if (kMask == 2) {
#pragma unroll
          for (int k = 0; k < 8; ++k) fa[k] = (vb[u] >> k) & 1u ? fa[k] : 0.f;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          s0[k] += fa[k];
          s1[k] = fmaf(fa[k], fx[k], s1[k]);
        }
      }
    }
  }
}

// Folds the per-thread accumulators over ty in shared memory; thread (ty==0) ends with the block total.
__device__ __forceinline__ void block_fold(float (&s0)[8], float (&s1)[8], int tx, int ty, int cx, int ry,
                                           float* smem /* [ry][cx][16] */) {
  float* mine = smem + ((size_t)ty * cx + tx) * 16;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    mine[k] = s0[k];
    mine[8 + k] = s1[k];
  }
  __syncthreads();
  if (ty == 0) {
    for (int j = 1; j < ry; ++j) {
      const float* o = smem + ((size_t)j * cx + tx) * 16;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        s0[k] += o[k];
        s1[k] += o[8 + k];
      }
    }
  }
  __syncthreads();
}

struct StatsScratch {      // layout of the caller-provided scratch buffer of pdb_bn_stats
  unsigned int ticket;     // blocks finished so far; the last one folds and resets it
  unsigned int pad[63];    // keep the partials 256-byte aligned
  // float partial[gridDim.x][2*C] follows
};

template <typename T>
__global__ void __launch_bounds__(kThreads) bn_stats_kernel(const T* __restrict__ x, int64_t rows, int C, int cx,
                                                            int ry, float* __restrict__ sums,
                                                            StatsScratch* __restrict__ scratch) {
  extern __shared__ float smem[];
  __shared__ bool is_last;
  const int tx = threadIdx.x % cx, ty = threadIdx.x / cx;
  float* partial = reinterpret_cast<float*>(scratch + 1) + (size_t)blockIdx.x * 2 * C;
  const int ncg = C / 8;
  for (int cg0 = 0; cg0 < ncg; cg0 += cx) {        // uniform trip count: block_fold contains barriers
    const int cg = cg0 + tx;
    float s0[8] = {0}, s1[8] = {0};
    if (ty < ry && cg < ncg) reduce_rows<T, 0, 0>(x, nullptr, nullptr, nullptr, rows, C, cg, cx, ry, ty, nullptr, nullptr, s0, s1);
    block_fold(s0, s1, tx, ty, cx, ry, smem);
    if (ty == 0 && cg < ncg) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        partial[cg * 8 + k] = s0[k];
        partial[C + cg * 8 + k] = s1[k];
      }
    }
  }
  // publish the partial, take a ticket; the last block folds all partials in block order
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(&scratch->ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (is_last) {
    __threadfence();
    const float* all = reinterpret_cast<const float*>(scratch + 1);
    for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) {
      float acc = 0.f;
      for (unsigned b = 0; b < gridDim.x; ++b) acc += __ldcg(all + (size_t)b * 2 * C + c);
      sums[c] = acc;
    }
    if (threadIdx.x == 0) scratch->ticket = 0;     // ready for the next launch on this scratch
  }
}

// K6: SyncBatchNorm statistics with the cross-GPU reduction INSIDE the statistics kernel.  The delegate runs
// batch_norm_stats -> all_gather (NCCL) -> batch_norm_gather_stats_with_counts, three launches and a library
// collective per BN layer and direction (torch:nn/modules/_functions.py:7-124).  Here the block that folds
// the per-block partials also pushes the [2C] sums into slot `rank` of every peer's symmetric staging
// buffer (P2P stores over NVLink), meets the peers on the signal pad, and adds up the `world` slots — the
// kernel returns global sums, no collective call.
template <typename T>
__global__ void __launch_bounds__(kThreads) bn_stats_sync_kernel(const T* __restrict__ x, int64_t rows, int C, int cx,
                                                                 int ry, float* __restrict__ sums,
                                                                 StatsScratch* __restrict__ scratch, SyncPeers sp) {
  extern __shared__ float smem[];
  __shared__ bool is_last;
  const int tx = threadIdx.x % cx, ty = threadIdx.x / cx;
  float* partial = reinterpret_cast<float*>(scratch + 1) + (size_t)blockIdx.x * 2 * C;
  const int ncg = C / 8;
  for (int cg0 = 0; cg0 < ncg; cg0 += cx) {        // uniform trip count: block_fold contains barriers
    const int cg = cg0 + tx;
    float s0[8] = {0}, s1[8] = {0};
    if (ty < ry && cg < ncg) reduce_rows<T, 0, 0>(x, nullptr, nullptr, nullptr, rows, C, cg, cx, ry, ty, nullptr, nullptr, s0, s1);
    block_fold(s0, s1, tx, ty, cx, ry, smem);
    if (ty == 0 && cg < ncg) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        partial[cg * 8 + k] = s0[k];
        partial[C + cg * 8 + k] = s1[k];
      }
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(&scratch->ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const float* all = reinterpret_cast<const float*>(scratch + 1);
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) {
    float acc = 0.f;
    for (unsigned b = 0; b < gridDim.x; ++b) acc += __ldcg(all + (size_t)b * 2 * C + c);
    sums[c] = acc;                                       // local sums; replaced by the global ones below
  }
  if (threadIdx.x == 0) scratch->ticket = 0;
  __syncthreads();
  exchange_sum(sp, sums, sums, 2 * C, threadIdx.x, blockDim.x, [] { __syncthreads(); });
}

// Every block ends by adding its per-channel partials to global memory.  Hundreds of blocks hitting the same 2C
// addresses serialise in the L2 atomic units (~100 ns per atomic per address: 45 us per launch with 444 blocks,
// more than the data movement of the small layers).  With `rep` the adds are spread over kReduceReplicas copies
// (block b -> copy b % R), and the last block to finish (ticket) folds the copies into dsums.
constexpr int kReduceReplicas = 16;

template <typename T, int kMask>
__global__ void __launch_bounds__(kThreads, PDB_BN_REDUCE_MINBLOCKS) bn_bwd_reduce_kernel(const T* __restrict__ dz, const T* __restrict__ z,
                                                                    const uint8_t* __restrict__ bits,
                                                                    const T* __restrict__ x, int64_t rows, int C, int cx,
                                                                    int ry, const float* __restrict__ mean,
                                                                    const float* __restrict__ rstd,
                                                                    float* __restrict__ dsums, float* __restrict__ rep,
                                                                    unsigned int* __restrict__ ticket, const SyncPeers sp) {
  extern __shared__ float smem[];
  __shared__ bool is_last;
  const int tx = threadIdx.x % cx, ty = threadIdx.x / cx;
  const int ncg = C / 8;
  float* target = rep ? rep + (size_t)(blockIdx.x % kReduceReplicas) * 2 * C : dsums;
  for (int cg0 = 0; cg0 < ncg; cg0 += cx) {        // uniform trip count: block_fold contains barriers
    const int cg = cg0 + tx;
    float s0[8] = {0}, s1[8] = {0};
    if (ty < ry && cg < ncg) reduce_rows<T, 1, kMask>(dz, z, bits, x, rows, C, cg, cx, ry, ty, mean, rstd, s0, s1);
    block_fold(s0, s1, tx, ty, cx, ry, smem);
    if (ty == 0 && cg < ncg) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int c = cg * 8 + k;
        atomicAdd(target + c, s0[k]);
        atomicAdd(target + C + c, rstd[c] * (s1[k] - mean[c] * s0[k]));     // block-local sum(g * xhat)
      }
    }
  }
  if (rep == nullptr) return;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (((reinterpret_cast<uintptr_t>(dsums) | reinterpret_cast<uintptr_t>(rep)) & 15) == 0) {      // C % 8 == 0 always
    for (int c4 = threadIdx.x; c4 < C / 2; c4 += blockDim.x) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < kReduceReplicas; ++r) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(rep + (size_t)r * 2 * C) + c4);
        a.x += v.x, a.y += v.y, a.z += v.z, a.w += v.w;
      }
      reinterpret_cast<float4*>(dsums)[c4] = a;
    }
  } else {
    for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) {
      float acc = 0.f;
#pragma unroll
      for (int r = 0; r < kReduceReplicas; ++r) acc += __ldcg(rep + (size_t)r * 2 * C + c);
      dsums[c] = acc;
    }
  }
  // SyncBatchNorm backward (torch:nn/modules/_functions.py:126-209 all-reduces [sum_dy, sum_dy_xmu]): the block that
  // folded the replicas exchanges the [2C] sums with its peers right here — no extra launch, one cross-GPU meeting
  if (sp.world > 1) {
    __syncthreads();
    exchange_sum(sp, dsums, dsums, 2 * C, threadIdx.x, blockDim.x, [] { __syncthreads(); });
  }
}

// Deterministic variant (PDB200_DETERMINISTIC=1): no atomics — every block writes its partials to a slot of its own
// and the last block adds the slots in block order, like bn_stats_kernel.  Two runs on the same input give the same
// bits, which turns "run it twice and compare" into a race detector for the whole step.
template <typename T, int kMask>
__global__ void __launch_bounds__(kThreads, 3) bn_bwd_reduce_det_kernel(const T* __restrict__ dz, const T* __restrict__ z,
                                                                        const uint8_t* __restrict__ bits,
                                                                        const T* __restrict__ x, int64_t rows, int C,
                                                                        int cx, int ry, const float* __restrict__ mean,
                                                                        const float* __restrict__ rstd,
                                                                        float* __restrict__ dsums,
                                                                        StatsScratch* __restrict__ scratch) {
  extern __shared__ float smem[];
  __shared__ bool is_last;
  const int tx = threadIdx.x % cx, ty = threadIdx.x / cx;
  float* partial = reinterpret_cast<float*>(scratch + 1) + (size_t)blockIdx.x * 2 * C;
  const int ncg = C / 8;
  for (int cg0 = 0; cg0 < ncg; cg0 += cx) {
    const int cg = cg0 + tx;
    float s0[8] = {0}, s1[8] = {0};
    if (ty < ry && cg < ncg) reduce_rows<T, 1, kMask>(dz, z, bits, x, rows, C, cg, cx, ry, ty, mean, rstd, s0, s1);
    block_fold(s0, s1, tx, ty, cx, ry, smem);
    if (ty == 0 && cg < ncg) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int c = cg * 8 + k;
        partial[c] = s0[k];
        partial[C + c] = rstd[c] * (s1[k] - mean[c] * s0[k]);
      }
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(&scratch->ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const float* all = reinterpret_cast<const float*>(scratch + 1);
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) {
    float acc = 0.f;
    for (unsigned b = 0; b < gridDim.x; ++b) acc += __ldcg(all + (size_t)b * 2 * C + c);
    dsums[c] = acc;
  }
  if (threadIdx.x == 0) scratch->ticket = 0;
}

// ------------------------------------------------------------------------------------------------
// Forward apply.
template <typename T, bool kRes, bool kRelu>
__global__ void __launch_bounds__(kThreads) bn_apply_kernel(const T* __restrict__ x, const T* __restrict__ res,
                                                            T* __restrict__ y, int64_t nvec, int C,
                                                            const float* __restrict__ sums, double count,
                                                            const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, float eps, float momentum,
                                                            float* running_mean, float* running_var,
                                                            int64_t* num_batches_tracked, float* save_mean,
                                                            float* save_rstd) {
  extern __shared__ float smem[];        // scale[C] | shift[C]
  float* scale = smem;
  float* shift = smem + C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const double m = (double)sums[c] / count;
    double v = (double)sums[C + c] / count - m * m;
    v = v > 0.0 ? v : 0.0;
    const float rstd = rsqrtf((float)v + eps);
    const float sc = gamma[c] * rstd;
    scale[c] = sc;
    shift[c] = beta[c] - (float)m * sc;
    if (blockIdx.x == 0) {
      if (save_mean) save_mean[c] = (float)m;
      if (save_rstd) save_rstd[c] = rstd;
      if (running_mean) {
        const double unbiased = count > 1.0 ? v * count / (count - 1.0) : v;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)m;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && num_batches_tracked) *num_batches_tracked += 1;
  __syncthreads();
  const int cvec = C / 8;
  const bool fixed_c = (blockDim.x % cvec) == 0;
  const int c_fixed = (int)(threadIdx.x % cvec) * 8;
  constexpr int U = 4;                     // independent 128-bit requests in flight per thread and stream
  const int64_t tile = (int64_t)blockDim.x * U;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < nvec; base += (int64_t)gridDim.x * tile) {
    V8<T> vx[U], vr[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
      if (i < nvec) {
        vx[u] = ld8s<T>(x + i * 8);
        if (kRes) vr[u] = ld8s<T>(res + i * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
      if (i >= nvec) continue;
      // 64-bit modulo is ~50 instructions; every base is a multiple of blockDim, so when cvec divides
      // blockDim (all power-of-two channel counts) the channel offset is a per-thread constant
      const int c0 = fixed_c ? c_fixed : (int)(i % cvec) * 8;
      float f[8];
      unpack8<T>(vx[u], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], scale[c0 + k], shift[c0 + k]);
      if (kRes) {
        float r[8];
        unpack8<T>(vr[u], r);
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] += r[k];
      }
      if (kRelu) {
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
      }
      st8<T>(y + i * 8, pack8v<T>(f));
    }
  }
}

template <typename T, bool kRes, bool kRelu>
__global__ void __launch_bounds__(kThreads) bn_apply_eval_kernel(const T* __restrict__ x, const T* __restrict__ res,
                                                                 T* __restrict__ y, int64_t nvec, int C,
                                                                 const float* __restrict__ rm,
                                                                 const float* __restrict__ rv,
                                                                 const float* __restrict__ gamma,
                                                                 const float* __restrict__ beta, float eps) {
  extern __shared__ float smem[];
  float* scale = smem;
  float* shift = smem + C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float sc = gamma[c] * rsqrtf(rv[c] + eps);
    scale[c] = sc;
    shift[c] = beta[c] - rm[c] * sc;
  }
  __syncthreads();
  const int cvec = C / 8;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const int c0 = (int)(i % cvec) * 8;
    float f[8];
    unpack8<T>(ld8s<T>(x + i * 8), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], scale[c0 + k], shift[c0 + k]);
    if (kRes) {
      float r[8];
      unpack8<T>(ld8s<T>(res + i * 8), r);
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] += r[k];
    }
    if (kRelu) {
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
    }
    st8<T>(y + i * 8, pack8v<T>(f));
  }
}

// ------------------------------------------------------------------------------------------------
// Backward apply: dx = A*g + B*x + Cc with per-channel A = gamma*rstd, B = -A*k2*rstd,
// Cc = -B*mean - A*k1, k1 = dsums[c]/count, k2 = dsums[C+c]/count.
template <typename T, bool kHasZ, bool kDres>
__global__ void __launch_bounds__(kThreads) bn_bwd_apply_kernel(const T* __restrict__ dz, const T* __restrict__ z,
                                                                const T* __restrict__ x, T* __restrict__ dx,
                                                                T* __restrict__ dres, int64_t nvec, int C,
                                                                const float* __restrict__ mean,
                                                                const float* __restrict__ rstd,
                                                                const float* __restrict__ gamma,
                                                                const float* __restrict__ dsums, double count,
                                                                float* dgamma, float* dbeta) {
  extern __shared__ float smem[];        // A[C] | B[C] | Cc[C]
  float* A = smem;
  float* B = smem + C;
  float* Cc = smem + 2 * C;
  const float inv = (float)(1.0 / count);
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float a = gamma[c] * rstd[c];
    const float k1 = dsums[c] * inv, k2 = dsums[C + c] * inv;
    const float b = -a * k2 * rstd[c];
    A[c] = a;
    B[c] = b;
    Cc[c] = -b * mean[c] - a * k1;
    if (blockIdx.x == 0) {
      if (dgamma) dgamma[c] += dsums[C + c];
      if (dbeta) dbeta[c] += dsums[c];
    }
  }
  __syncthreads();
  const int cvec = C / 8;
  const bool fixed_c = (blockDim.x % cvec) == 0;
  const int c_fixed = (int)(threadIdx.x % cvec) * 8;
  constexpr int U = 2;
  const int64_t tile = (int64_t)blockDim.x * U;
  for (int64_t base = (int64_t)blockIdx.x * tile; base < nvec; base += (int64_t)gridDim.x * tile) {
    V8<T> vg[U], vx[U], vz[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
      if (i < nvec) {
        vg[u] = ld8s<T>(dz + i * 8);
        vx[u] = ld8s<T>(x + i * 8);
        if (kHasZ) vz[u] = ld8s<T>(z + i * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = base + (int64_t)u * blockDim.x + threadIdx.x;
      if (i >= nvec) continue;
      const int c0 = fixed_c ? c_fixed : (int)(i % cvec) * 8;
      float g[8], fx[8];
      unpack8<T>(vg[u], g);
      unpack8<T>(vx[u], fx);
      if (kHasZ) {
        float fz[8];
        unpack8<T>(vz[u], fz);
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] = fz[k] > 0.f ? g[k] : 0.f;
      }
      if (kDres) st8<T>(dres + i * 8, pack8v<T>(g));
      float o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = fmaf(A[c0 + k], g[k], fmaf(B[c0 + k], fx[k], Cc[c0 + k]));
      st8<T>(dx + i * 8, pack8v<T>(o));
    }
  }
}

static void reduce_geometry(int C, int* cx, int* ry) {
  int ncg = C / 8;
  int x = ncg < kThreads ? ncg : kThreads;
  *cx = x;
  *ry = kThreads / x;
}

// Persistent grid-stride kernels must be launched as exactly ONE wave: with more blocks than can be resident the
// leftover blocks run alone on a mostly idle GPU (3 blocks/SM resident and a grid of 4/SM = 1.33 waves cost 4 %).
template <typename K>
static int resident_blocks(K kernel, size_t smem) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kThreads, smem) != cudaSuccess || per_sm < 1) per_sm = 2;
  return per_sm * num_sms();
}

static int reduce_grid(int64_t rows, int ry, int cap = 0) {
  int64_t want = (rows + (int64_t)ry * 4 - 1) / ((int64_t)ry * 4);      // a block walks slabs of 4*ry rows
  if (cap <= 0) cap = num_sms() * 4;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

static int ew_grid(int64_t nvec) {
  int64_t want = (nvec + kThreads - 1) / kThreads;
  int64_t cap = (int64_t)num_sms() * 8;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace pdb

using namespace pdb;

extern "C" size_t pdb_bn_stats_scratch_bytes(int C) {
  return sizeof(StatsScratch) + (size_t)num_sms() * 4 * 2 * (size_t)C * sizeof(float);
}

extern "C" int pdb_bn_stats(const void* x, int dtype, int64_t rows, int C, float* sums, void* scratch,
                            size_t scratch_bytes, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_bn_stats", x);
  PDB_CHECK_ARG(x && sums && scratch, "pdb_bn_stats: null pointer");
  PDB_CHECK_ARG(C > 0 && C % 8 == 0 && C <= kMaxC, "pdb_bn_stats: C=%d must be a multiple of 8 and <= %d", C, kMaxC);
  PDB_CHECK_ARG(rows > 0, "pdb_bn_stats: rows must be positive");
  int cx, ry;
  reduce_geometry(C, &cx, &ry);
  const int grid = reduce_grid(rows, ry);
  const size_t need = sizeof(StatsScratch) + (size_t)grid * 2 * C * sizeof(float);
  PDB_CHECK_ARG(scratch_bytes >= need, "pdb_bn_stats: scratch too small (%zu < %zu bytes)", scratch_bytes, need);
  PDB_CHECK_ARG(((uintptr_t)scratch) % 256 == 0, "pdb_bn_stats: scratch must be 256-byte aligned");
  const size_t smem = (size_t)kThreads * 16 * sizeof(float);
  PDB_DISPATCH_ACT(dtype, T, (bn_stats_kernel<T><<<grid, kThreads, smem, (cudaStream_t)stream>>>(
                                (const T*)x, rows, C, cx, ry, sums, (StatsScratch*)scratch)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_bn_stats_sync(const void* x, int dtype, int64_t rows, int C, float* sums, void* scratch,
                                 size_t scratch_bytes, void* const* staging, void* const* pads, int rank, int world,
                                 int slot_stride, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_bn_stats_sync", x);
  PDB_CHECK_ARG(x && sums && scratch && staging && pads, "pdb_bn_stats_sync: null pointer");
  PDB_CHECK_ARG(C > 0 && C % 8 == 0 && C <= kMaxC && rows > 0, "pdb_bn_stats_sync: bad shape rows=%lld C=%d", (long long)rows, C);
  PDB_CHECK_ARG(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, "pdb_bn_stats_sync: bad rank/world");
  PDB_CHECK_ARG(2 * C <= slot_stride, "pdb_bn_stats_sync: 2C=%d exceeds the staging slot (%d floats)", 2 * C, slot_stride);
  int cx, ry;
  reduce_geometry(C, &cx, &ry);
  const int grid = reduce_grid(rows, ry);
  const size_t need = sizeof(StatsScratch) + (size_t)grid * 2 * C * sizeof(float);
  PDB_CHECK_ARG(scratch_bytes >= need && ((uintptr_t)scratch) % 256 == 0, "pdb_bn_stats_sync: scratch too small or misaligned");
  SyncPeers sp;
  if (int e = fill_sync(&sp, staging, pads, rank, world, slot_stride, 0)) return e;
  const size_t smem = (size_t)kThreads * 16 * sizeof(float);
  PDB_DISPATCH_ACT(dtype, T, (bn_stats_sync_kernel<T><<<grid, kThreads, smem, (cudaStream_t)stream>>>(
                                (const T*)x, rows, C, cx, ry, sums, (StatsScratch*)scratch, sp)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" size_t pdb_bn_bwd_reduce_scratch_floats(int C) { return (size_t)kReduceReplicas * 2 * C + 32; }

extern "C" int pdb_bn_bwd_reduce(const void* dz, const void* z, const uint8_t* relu_mask, const void* x, int dtype,
                                 int64_t rows, int C, const float* mean, const float* rstd, float* dsums,
                                 float* scratch, const pdb_sync_peers* sync, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_bn_bwd_reduce", dz, z, x);
  PDB_CHECK_ARG(dz && x && mean && rstd && dsums, "pdb_bn_bwd_reduce: null pointer");
  SyncPeers sp{};
  if (sync && sync->world > 1) {
    PDB_CHECK_ARG(scratch, "pdb_bn_bwd_reduce: the cross-GPU exchange needs the replica scratch (last-block fold)");
    PDB_CHECK_ARG(2 * C <= sync->slot_stride, "pdb_bn_bwd_reduce: 2C=%d exceeds the staging slot (%d floats)", 2 * C, sync->slot_stride);
    if (int e = fill_sync(&sp, sync->staging, sync->pads, sync->rank, sync->world, sync->slot_stride, sync->row)) return e;
  }
  float* rep = scratch;
  unsigned int* ticket = scratch ? reinterpret_cast<unsigned int*>(scratch + (size_t)kReduceReplicas * 2 * C) : nullptr;
  PDB_CHECK_ARG(C > 0 && C % 8 == 0 && C <= kMaxC && rows > 0, "pdb_bn_bwd_reduce: bad shape rows=%lld C=%d",
                (long long)rows, C);
  int cx, ry;
  reduce_geometry(C, &cx, &ry);
  const size_t smem = (size_t)kThreads * 16 * sizeof(float);
  static int cap = 0;
  if (cap == 0) cap = resident_blocks(bn_bwd_reduce_kernel<__nv_bfloat16, 2>, smem);
  const int grid = reduce_grid(rows, ry, cap);
  cudaStream_t st = (cudaStream_t)stream;
  PDB_DISPATCH_ACT(dtype, T, {
    if (relu_mask)
      bn_bwd_reduce_kernel<T, 2><<<grid, kThreads, smem, st>>>((const T*)dz, nullptr, relu_mask, (const T*)x, rows, C,
                                                               cx, ry, mean, rstd, dsums, rep, ticket, sp);
    else if (z)
      bn_bwd_reduce_kernel<T, 1><<<grid, kThreads, smem, st>>>((const T*)dz, (const T*)z, nullptr, (const T*)x, rows, C,
                                                               cx, ry, mean, rstd, dsums, rep, ticket, sp);
    else
      bn_bwd_reduce_kernel<T, 0><<<grid, kThreads, smem, st>>>((const T*)dz, nullptr, nullptr, (const T*)x, rows, C, cx,
                                                               ry, mean, rstd, dsums, rep, ticket, sp);
  });
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_bn_bwd_reduce_det(const void* dz, const void* z, const uint8_t* relu_mask, const void* x, int dtype,
                                     int64_t rows, int C, const float* mean, const float* rstd, float* dsums,
                                     void* scratch, size_t scratch_bytes, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_bn_bwd_reduce_det", dz, z, x);
  PDB_CHECK_ARG(dz && x && mean && rstd && dsums && scratch, "pdb_bn_bwd_reduce_det: null pointer");
  PDB_CHECK_ARG(C > 0 && C % 8 == 0 && C <= kMaxC && rows > 0, "pdb_bn_bwd_reduce_det: bad shape rows=%lld C=%d",
                (long long)rows, C);
  int cx, ry;
  reduce_geometry(C, &cx, &ry);
  const size_t smem = (size_t)kThreads * 16 * sizeof(float);
  static int cap = 0;
  if (cap == 0) cap = resident_blocks(bn_bwd_reduce_det_kernel<__nv_bfloat16, 2>, smem);
  const int grid = reduce_grid(rows, ry, cap);
  const size_t need = sizeof(StatsScratch) + (size_t)grid * 2 * C * sizeof(float);
  PDB_CHECK_ARG(scratch_bytes >= need && ((uintptr_t)scratch) % 256 == 0,
                "pdb_bn_bwd_reduce_det: scratch too small (%zu < %zu bytes) or misaligned", scratch_bytes, need);
  cudaStream_t st = (cudaStream_t)stream;
  StatsScratch* sc = (StatsScratch*)scratch;
  PDB_DISPATCH_ACT(dtype, T, {
    if (relu_mask)
      bn_bwd_reduce_det_kernel<T, 2><<<grid, kThreads, smem, st>>>((const T*)dz, nullptr, relu_mask, (const T*)x, rows, C, cx, ry, mean, rstd, dsums, sc);
    else if (z)
      bn_bwd_reduce_det_kernel<T, 1><<<grid, kThreads, smem, st>>>((const T*)dz, (const T*)z, nullptr, (const T*)x, rows, C, cx, ry, mean, rstd, dsums, sc);
    else
      bn_bwd_reduce_det_kernel<T, 0><<<grid, kThreads, smem, st>>>((const T*)dz, nullptr, nullptr, (const T*)x, rows, C, cx, ry, mean, rstd, dsums, sc);
  });
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_bn_apply(const void* x, const void* residual, void* y, int dtype, int64_t rows, int C,
                            const float* sums, double count, const float* gamma, const float* beta, float eps,
                            float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked,
                            float* save_mean, float* save_rstd, int relu, uint8_t* relu_mask, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_bn_apply", x, residual, y);
  PDB_CHECK_ARG(x && y && sums && gamma && beta, "pdb_bn_apply: null pointer");
  PDB_CHECK_ARG(!relu_mask || (relu && bn_fast_ok(C)),
                "pdb_bn_apply: relu_mask needs relu and a channel count whose 8-channel groups divide 256 (C=%d)", C);
  PDB_CHECK_ARG(C > 0 && C % 8 == 0 && C <= kMaxC && rows > 0 && count > 0, "pdb_bn_apply: bad shape rows=%lld C=%d",
                (long long)rows, C);
  const int64_t nvec = rows * C / 8;
  cudaStream_t st = (cudaStream_t)stream;
  if (bn_fast_ok(C)) {
    const int fgrid = bn_fast_grid(nvec, kFastU, dtype);
    const float inv_count = (float)(1.0 / count), unbias = (float)(count > 1.0 ? count / (count - 1.0) : 1.0);
#define FAST(RES, RELU)                                                                                            \
  bn_apply_fast_kernel<T, RES, RELU><<<fgrid, kFastThreads, 0, st>>>(                                              \
      (const T*)x, (const T*)residual, (T*)y, nvec, C, sums, inv_count, unbias, gamma, beta, eps, momentum,        \
      running_mean, running_var, num_batches_tracked, save_mean, save_rstd, relu_mask)
    PDB_DISPATCH_ACT(dtype, T, {
      if (residual && relu) FAST(true, true);
      else if (residual) FAST(true, false);
      else if (relu) FAST(false, true);
      else FAST(false, false);
    });
#undef FAST
    PDB_LAUNCH_CHECK();
    return 0;
  }
  const int grid = ew_grid(nvec);
  const size_t smem = 2 * (size_t)C * sizeof(float);
#define LAUNCH(RES, RELU)                                                                                         \
  bn_apply_kernel<T, RES, RELU><<<grid, kThreads, smem, st>>>((const T*)x, (const T*)residual, (T*)y, nvec, C, sums, \
                                                              count, gamma, beta, eps, momentum, running_mean,    \
                                                              running_var, num_batches_tracked, save_mean, save_rstd)
  PDB_DISPATCH_ACT(dtype, T, {
    if (residual && relu) LAUNCH(true, true);
    else if (residual) LAUNCH(true, false);
    else if (relu) LAUNCH(false, true);
    else LAUNCH(false, false);
  });
#undef LAUNCH
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_bn_apply_eval(const void* x, const void* residual, void* y, int dtype, int64_t rows, int C,
                                 const float* running_mean, const float* running_var, const float* gamma,
                                 const float* beta, float eps, int relu, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_bn_apply_eval", x, residual, y);
  PDB_CHECK_ARG(x && y && running_mean && running_var && gamma && beta, "pdb_bn_apply_eval: null pointer");
  PDB_CHECK_ARG(C > 0 && C % 8 == 0 && C <= kMaxC && rows > 0, "pdb_bn_apply_eval: bad shape");
  const int64_t nvec = rows * C / 8;
  cudaStream_t st = (cudaStream_t)stream;
  if (bn_fast_ok(C)) {
    const int fgrid = bn_fast_grid(nvec, kFastU, dtype);
#define FAST(RES, RELU)                                                                                     \
  bn_apply_eval_fast_kernel<T, RES, RELU><<<fgrid, kFastThreads, 0, st>>>(                                  \
      (const T*)x, (const T*)residual, (T*)y, nvec, C, running_mean, running_var, gamma, beta, eps)
    PDB_DISPATCH_ACT(dtype, T, {
      if (residual && relu) FAST(true, true);
      else if (residual) FAST(true, false);
      else if (relu) FAST(false, true);
      else FAST(false, false);
    });
#undef FAST
    PDB_LAUNCH_CHECK();
    return 0;
  }
  const int grid = ew_grid(nvec);
  const size_t smem = 2 * (size_t)C * sizeof(float);
#define LAUNCH(RES, RELU)                                                                                    \
  bn_apply_eval_kernel<T, RES, RELU><<<grid, kThreads, smem, st>>>((const T*)x, (const T*)residual, (T*)y, nvec, C, \
                                                                   running_mean, running_var, gamma, beta, eps)
  PDB_DISPATCH_ACT(dtype, T, {
    if (residual && relu) LAUNCH(true, true);
    else if (residual) LAUNCH(true, false);
    else if (relu) LAUNCH(false, true);
    else LAUNCH(false, false);
  });
#undef LAUNCH
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_bn_bwd_apply(const void* dz, const void* z, const uint8_t* relu_mask, const void* x, void* dx,
                                void* dres, int dtype, int64_t rows, int C, const float* mean, const float* rstd,
                                const float* gamma, const float* dsums, double count, float* dgamma, float* dbeta,
                                void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_bn_bwd_apply", dz, z, x, dx, dres);
  PDB_CHECK_ARG(dz && x && dx && mean && rstd && gamma && dsums, "pdb_bn_bwd_apply: null pointer");
  PDB_CHECK_ARG(!relu_mask || bn_fast_ok(C), "pdb_bn_bwd_apply: relu_mask needs a power-of-two channel count (C=%d)", C);
  PDB_CHECK_ARG(C > 0 && C % 8 == 0 && C <= kMaxC && rows > 0 && count > 0, "pdb_bn_bwd_apply: bad shape");
  const int64_t nvec = rows * C / 8;
  cudaStream_t st = (cudaStream_t)stream;
  if (bn_fast_ok(C)) {
    const int fgrid = bn_fast_grid(nvec, kFastUBwd, dtype);
    const float inv_count = (float)(1.0 / count);
#define FAST(MK, DR)                                                                                              \
  bn_bwd_apply_fast_kernel<T, MK, DR><<<fgrid, kFastThreads, 0, st>>>(                                            \
      (const T*)dz, (const T*)z, relu_mask, (const T*)x, (T*)dx, (T*)dres, nvec, C, mean, rstd, gamma, dsums,     \
      inv_count, dgamma, dbeta)
    PDB_DISPATCH_ACT(dtype, T, {
      if (relu_mask && dres) FAST(2, true);
      else if (relu_mask) FAST(2, false);
      else if (z && dres) FAST(1, true);
      else if (z) FAST(1, false);
      else if (dres) FAST(0, true);
      else FAST(0, false);
    });
#undef FAST
    PDB_LAUNCH_CHECK();
    return 0;
  }
  const int grid = ew_grid(nvec);
  const size_t smem = 3 * (size_t)C * sizeof(float);
#define LAUNCH(HZ, DR)                                                                                              \
  bn_bwd_apply_kernel<T, HZ, DR><<<grid, kThreads, smem, st>>>((const T*)dz, (const T*)z, (const T*)x, (T*)dx,     \
                                                               (T*)dres, nvec, C, mean, rstd, gamma, dsums, count, \
                                                               dgamma, dbeta)
  PDB_DISPATCH_ACT(dtype, T, {
    if (z && dres) LAUNCH(true, true);
    else if (z) LAUNCH(true, false);
    else if (dres) LAUNCH(false, true);
    else LAUNCH(false, false);
  });
#undef LAUNCH
  PDB_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// pdb_bn_stats_noalloc — the same statistics kernel for callers that cannot allocate device memory.
//
// pdb_bn_stats needs a caller-provided scratch buffer for the per-block partial sums.  Some users of
// this library have no way to produce one at the point of the call: a C/C++ host that gave all of
// the GPU's memory to its own pre-planned pools at start-up and has no allocator left to ask; code
// that runs in the middle of a training step, where cudaMalloc/cudaFree would synchronise the whole
// device and break the overlap between the compute and communication streams; embedders that are
// simply not allowed to call the CUDA allocator (the memory budget is owned by somebody else).
// For them the scratch lives here, in a fixed array in the library's device image: every GPU that
// loads libpdb200.so carries these kNoAllocScratchBytes (~512 KB) until the process exits, whether
// or not this entry point is ever called.
//
// Limits and rules, all consequences of the fixed array:
//   * C <= kNoAllocMaxC (512); larger layers are rejected with an error — use pdb_bn_stats;
//   * at most kNoAllocMaxBlocks (128) CTAs fold partials, whatever `rows` is;
//   * the array is shared by every call on the device and still holds the previous call's partials:
//     the block ticket is cleared on the stream before each launch, partials are fully overwritten
//     before they are read; two calls that run concurrently on one GPU (different streams) race on
//     it and produce garbage — serialise them;
//   * the first call on each device resolves the array's device address (cudaGetSymbolAddress; the
//     name of a __device__ array is only a host-side handle, not something a kernel can
//     dereference) — make that first call outside of any stream capture.
// Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
namespace pdb {
constexpr int kNoAllocMaxC = 512;
constexpr int kNoAllocMaxBlocks = 128;
constexpr size_t kNoAllocScratchBytes =
    sizeof(StatsScratch) + (size_t)kNoAllocMaxBlocks * 2 * kNoAllocMaxC * sizeof(float);
static __device__ __align__(256) unsigned char g_bn_stats_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
static void* g_noalloc_ptr[kMaxDevices] = {};      // per-device address of g_bn_stats_scratch
}  // namespace pdb

extern "C" int pdb_bn_stats_noalloc(const void* x, int dtype, int64_t rows, int C, float* sums, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_bn_stats_noalloc", x);
  PDB_CHECK_ARG(x && sums, "pdb_bn_stats_noalloc: null pointer");
  PDB_CHECK_ARG(C > 0 && C % 8 == 0, "pdb_bn_stats_noalloc: C=%d must be a positive multiple of 8", C);
  PDB_CHECK_ARG(C <= kNoAllocMaxC,
                "pdb_bn_stats_noalloc: C=%d exceeds the built-in scratch limit of %d channels; "
                "use pdb_bn_stats with a caller-provided scratch buffer", C, kNoAllocMaxC);
  PDB_CHECK_ARG(rows > 0, "pdb_bn_stats_noalloc: rows must be positive");
  int dev = -1;
  PDB_CUDA(cudaGetDevice(&dev));
  PDB_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "pdb_bn_stats_noalloc: device index %d out of range", dev);
  if (g_noalloc_ptr[dev] == nullptr) {
    // one-time set-up for this device (it is the current one): symbol -> device address
    void* p = nullptr;
    PDB_CUDA(cudaGetSymbolAddress(&p, g_bn_stats_scratch));
    PDB_CHECK_ARG(p != nullptr, "pdb_bn_stats_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[dev] = p;
  }
  StatsScratch* scratch = static_cast<StatsScratch*>(g_noalloc_ptr[dev]);
  int cx, ry;
  reduce_geometry(C, &cx, &ry);
  int grid = reduce_grid(rows, ry);
  if (grid > kNoAllocMaxBlocks) grid = kNoAllocMaxBlocks;
  cudaStream_t st = (cudaStream_t)stream;
  // The ticket counter is the only state that must start at zero; an earlier launch that was
  // aborted (or a racing one) could have left it non-zero.
  PDB_CUDA(cudaMemsetAsync(scratch, 0, sizeof(StatsScratch), st));
  const size_t smem = (size_t)kThreads * 16 * sizeof(float);
  PDB_DISPATCH_ACT(dtype, T, (bn_stats_kernel<T><<<grid, kThreads, smem, st>>>((const T*)x, rows, C, cx, ry, sums, scratch)));
  PDB_LAUNCH_CHECK();
  return 0;
}
