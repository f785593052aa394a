// Fast paths of the BatchNorm apply kernels for channel counts whose 8-channel groups divide the block
// size (every power-of-two C <= 2048, i.e. all of ResNet).  Measured on B200 (bench_tools/bw_probe.cu,
// profiles/bw_probe_r1.log): for a 2-read + 1-write 16-bit stream a "one tile per block" launch reaches
// 6.9 TB/s where a grid-stride loop stalls at 6.3 TB/s, and the first version of these kernels — per-channel
// scale/shift in shared memory, read with an 8-way bank conflict (lanes 32 B apart) — only got 4.3 TB/s.
// Here every thread owns the same 8 channels for its whole life, so the per-channel constants live in
// registers: they are derived once per thread from the raw sums (a handful of L2-resident 128-bit loads),
// there is no shared memory, no __syncthreads and no integer modulo in the loop.
#pragma once
#include "common.cuh"

namespace pdb {

constexpr int kFastThreads = 256;
constexpr int kFastU = 8;                 // 128-bit vectors per thread and tile
// fp32 activations are two 128-bit vectors per 8 channels: half as many 8-channel groups per thread keeps the bytes
// (and registers) in flight the same
template <typename T>
constexpr int fast_u(int u16) { return sizeof(T) == 4 ? u16 / 2 : u16; }

__device__ __forceinline__ void load8f(const float* __restrict__ p, float (&v)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w, v[4] = b.x, v[5] = b.y, v[6] = b.z, v[7] = b.w;
}

__device__ __forceinline__ void bn_moments(float s, float ss, float inv_count, float eps, float* mean, float* var,
                                           float* rstd) {
  const float m = s * inv_count;
  const float v = fmaxf(fmaf(ss, inv_count, -m * m), 0.f);
  *mean = m, *var = v, *rstd = rsqrtf(v + eps);
}

template <typename T, bool kRes, bool kRelu>
__global__ void __launch_bounds__(kFastThreads) bn_apply_fast_kernel(
    const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y, int64_t nvec, int C,
    const float* __restrict__ sums, float inv_count, float unbias, const float* __restrict__ gamma,
    const float* __restrict__ beta, float eps, float momentum, float* running_mean, float* running_var,
    int64_t* num_batches_tracked, float* save_mean, float* save_rstd, uint8_t* __restrict__ relu_mask) {
  const int cvec = C >> 3;
  const int c0 = (int)(threadIdx.x % cvec) << 3;
  float sc[8], sh[8];
  {
    float s[8], ss[8], g[8], b[8];
    load8f(sums + c0, s), load8f(sums + C + c0, ss), load8f(gamma + c0, g), load8f(beta + c0, b);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float m, v, r;
      bn_moments(s[k], ss[k], inv_count, eps, &m, &v, &r);
      sc[k] = g[k] * r;
      sh[k] = fmaf(-m, sc[k], b[k]);
    }
  }
  if (blockIdx.x == 0) {          // per-layer bookkeeping, once: saved moments and running statistics
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      float m, v, r;
      bn_moments(sums[c], sums[C + c], inv_count, eps, &m, &v, &r);
      if (save_mean) save_mean[c] = m;
      if (save_rstd) save_rstd[c] = r;
      if (running_mean) {
        running_mean[c] = fmaf(momentum, m - running_mean[c], running_mean[c]);
        running_var[c] = fmaf(momentum, v * unbias - running_var[c], running_var[c]);
      }
    }
    if (threadIdx.x == 0 && num_batches_tracked) *num_batches_tracked += 1;
  }
  constexpr int U = fast_u<T>(kFastU);
  const int64_t base = (int64_t)blockIdx.x * (kFastThreads * U) + threadIdx.x;
  V8<T> vx[U], vr[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int64_t i = base + (int64_t)u * kFastThreads;
    if (i < nvec) {
      vx[u] = ld8s<T>(x + i * 8);
      if (kRes) vr[u] = ld8s<T>(res + i * 8);
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int64_t i = base + (int64_t)u * kFastThreads;
    if (i >= nvec) continue;
    float f[8];
    unpack8<T>(vx[u], f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], sh[k]);
    if (kRes) {
      float r[8];
      unpack8<T>(vr[u], r);
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] += r[k];
    }
    if (kRelu) {
      if (relu_mask) {       // one bit per element: the backward pass reads this instead of the 16x larger z
        uint32_t m = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) m |= (f[k] > 0.f ? 1u : 0u) << k;
        relu_mask[i] = (uint8_t)m;
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
    }
    st8<T>(y + i * 8, pack8v<T>(f));
  }
}

template <typename T, bool kRes, bool kRelu>
__global__ void __launch_bounds__(kFastThreads) bn_apply_eval_fast_kernel(
    const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y, int64_t nvec, int C,
    const float* __restrict__ rm, const float* __restrict__ rv, const float* __restrict__ gamma,
    const float* __restrict__ beta, float eps) {
  const int cvec = C >> 3;
  const int c0 = (int)(threadIdx.x % cvec) << 3;
  float sc[8], sh[8];
  {
    float m[8], v[8], g[8], b[8];
    load8f(rm + c0, m), load8f(rv + c0, v), load8f(gamma + c0, g), load8f(beta + c0, b);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      sc[k] = g[k] * rsqrtf(v[k] + eps);
      sh[k] = fmaf(-m[k], sc[k], b[k]);
    }
  }
  constexpr int U = fast_u<T>(kFastU);
  const int64_t base = (int64_t)blockIdx.x * (kFastThreads * U) + threadIdx.x;
  V8<T> vx[U], vr[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int64_t i = base + (int64_t)u * kFastThreads;
    if (i < nvec) {
      vx[u] = ld8s<T>(x + i * 8);
      if (kRes) vr[u] = ld8s<T>(res + i * 8);
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int64_t i = base + (int64_t)u * kFastThreads;
    if (i >= nvec) continue;
    float f[8];
    unpack8<T>(vx[u], f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], sh[k]);
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

// dx = A*g + B*x + Cc,  A = gamma*rstd, B = -A*k2*rstd, Cc = -B*mean - A*k1, k1 = dsums[c]/count, k2 = dsums[C+c]/count
constexpr int kFastUBwd = 4;      // three input streams: fewer vectors per thread keep the registers in check
template <typename T, int kMask /* 0 none, 1 z > 0, 2 bit mask */, bool kDres>
__global__ void __launch_bounds__(kFastThreads) bn_bwd_apply_fast_kernel(
    const T* __restrict__ dz, const T* __restrict__ z, const uint8_t* __restrict__ bits, const T* __restrict__ x,
    T* __restrict__ dx, T* __restrict__ dres,
    int64_t nvec, int C, const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
    const float* __restrict__ dsums, float inv_count, float* dgamma, float* dbeta) {
  const int cvec = C >> 3;
  const int c0 = (int)(threadIdx.x % cvec) << 3;
  float A[8], B[8], Cc[8];
  {
    float m[8], r[8], g[8], d0[8], d1[8];
    load8f(mean + c0, m), load8f(rstd + c0, r), load8f(gamma + c0, g), load8f(dsums + c0, d0), load8f(dsums + C + c0, d1);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float a = g[k] * r[k];
      const float b = -a * (d1[k] * inv_count) * r[k];
      A[k] = a, B[k] = b;
      Cc[k] = -b * m[k] - a * (d0[k] * inv_count);
    }
  }
  if (blockIdx.x == 0 && (dgamma || dbeta)) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      if (dgamma) dgamma[c] += dsums[C + c];
      if (dbeta) dbeta[c] += dsums[c];
    }
  }
  constexpr int U = fast_u<T>(kFastUBwd);
  const int64_t base = (int64_t)blockIdx.x * (kFastThreads * U) + threadIdx.x;
  V8<T> vg[U], vx[U], vz[U];
  uint32_t vb[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int64_t i = base + (int64_t)u * kFastThreads;
    if (i < nvec) {
      vg[u] = ld8s<T>(dz + i * 8);
      vx[u] = ld8s<T>(x + i * 8);
      if (kMask == 1) vz[u] = ld8s<T>(z + i * 8);
      if (kMask == 2) vb[u] = __ldg(bits + i);
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int64_t i = base + (int64_t)u * kFastThreads;
    if (i >= nvec) continue;
    float g[8], fx[8];
    unpack8<T>(vg[u], g);
    unpack8<T>(vx[u], fx);
    if (kMask == 1) {
      float fz[8];
      unpack8<T>(vz[u], fz);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = fz[k] > 0.f ? g[k] : 0.f;
    }
    if (kMask == 2) {
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = (vb[u] >> k) & 1u ? g[k] : 0.f;
    }
    if (kDres) st8<T>(dres + i * 8, pack8v<T>(g));
    float o[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) o[k] = fmaf(A[k], g[k], fmaf(B[k], fx[k], Cc[k]));
    st8<T>(dx + i * 8, pack8v<T>(o));
  }
}

inline bool bn_fast_ok(int C) { return C >= 8 && (C % 8) == 0 && (kFastThreads % (C / 8)) == 0; }
inline int bn_fast_grid(int64_t nvec, int u16, int dtype) {
  const int u = dtype == PDB_FP32 ? u16 / 2 : u16;
  return (int)((nvec + (int64_t)kFastThreads * u - 1) / ((int64_t)kFastThreads * u)); }

}  // namespace pdb
