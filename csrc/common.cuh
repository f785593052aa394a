// Shared helpers for the sm_100a kernel library.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>

#include "pdb200.h"

namespace pdb {

void set_error(const char* fmt, ...);   // defined in error.cu

#define PDB_CHECK_ARG(cond, ...)            \
  do {                                      \
    if (!(cond)) {                          \
      ::pdb::set_error(__VA_ARGS__);        \
      return 1;                             \
    }                                       \
  } while (0)

#define PDB_CUDA(expr)                                                                     \
  do {                                                                                     \
    cudaError_t e__ = (expr);                                                              \
    if (e__ != cudaSuccess) {                                                              \
      ::pdb::set_error("%s: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return 2;                                                                            \
    }                                                                                      \
  } while (0)

#define PDB_LAUNCH_CHECK() PDB_CUDA(cudaGetLastError())

inline int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// ---- 16-bit element traits --------------------------------------------------------------
template <typename T> struct Elem;
template <> struct Elem<__nv_bfloat16> {
  using T2 = __nv_bfloat162;
  static __device__ __forceinline__ float2 to_f2(T2 v) { return __bfloat1622float2(v); }
  static __device__ __forceinline__ T2 from_f2(float a, float b) { return __floats2bfloat162_rn(a, b); }
  static __device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
};
template <> struct Elem<__half> {
  using T2 = __half2;
  static __device__ __forceinline__ float2 to_f2(T2 v) { return __half22float2(v); }
  static __device__ __forceinline__ T2 from_f2(float a, float b) { return __floats2half2_rn(a, b); }
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
};

template <> struct Elem<float> {      // fp32 storage (TF32 tensor-core math): the "conversions" are identities
  using T2 = float2;
  static __device__ __forceinline__ float2 to_f2(T2 v) { return v; }
  static __device__ __forceinline__ T2 from_f2(float a, float b) { return make_float2(a, b); }
  static __device__ __forceinline__ float to_f(float v) { return v; }
  static __device__ __forceinline__ float from_f(float v) { return v; }
};

template <typename T>
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  using T2 = typename Elem<T>::T2;
  const T2* h = reinterpret_cast<const T2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = Elem<T>::to_f2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

template <typename T>
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  using T2 = typename Elem<T>::T2;
  uint4 v;
  T2* h = reinterpret_cast<T2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = Elem<T>::from_f2(f[2 * i], f[2 * i + 1]);
  return v;
}

__device__ __forceinline__ uint4 ldg128(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
// streaming load: read-once data should not displace reusable lines in L1
__device__ __forceinline__ uint4 ldg128_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg128(void* p, const uint4& v) { *reinterpret_cast<uint4*>(p) = v; }

// ---- 8 consecutive elements of any activation type: one 128-bit access (16-bit) or two (fp32) ----------------
template <typename T>
struct V8 {
  uint4 q[sizeof(T) / 2];
};
// fp32: ONE 256-bit access per lane (sm_100 has LDG.256 / STG.256).  Two 128-bit accesses per lane touch every 32-byte
// sector twice with half of it used each time; with the no-allocate hint the second touch goes back to L2, i.e. twice
// the L2->SM sectors of the 16-bit kernels for the same bytes (the fp32 BatchNorm passes ran at 5.4 TB/s where the
// 16-bit ones reach 6.3).  Needs 32-byte aligned addresses: the entry points check that for PDB_FP32.
template <typename T>
__device__ __forceinline__ V8<T> ld8(const T* p) {
  V8<T> v;
  if constexpr (sizeof(T) == 4) {
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v.q[0].x), "=r"(v.q[0].y), "=r"(v.q[0].z), "=r"(v.q[0].w), "=r"(v.q[1].x), "=r"(v.q[1].y), "=r"(v.q[1].z), "=r"(v.q[1].w)
                 : "l"(p));
  } else {
    v.q[0] = __ldg(reinterpret_cast<const uint4*>(p));
  }
  return v;
}
template <typename T>
__device__ __forceinline__ V8<T> ld8s(const T* p) {           // streaming (read-once) variant
  V8<T> v;
  if constexpr (sizeof(T) == 4) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v.q[0].x), "=r"(v.q[0].y), "=r"(v.q[0].z), "=r"(v.q[0].w), "=r"(v.q[1].x), "=r"(v.q[1].y), "=r"(v.q[1].z), "=r"(v.q[1].w)
                 : "l"(p));
  } else {
    v.q[0] = ldg128_stream(reinterpret_cast<const uint4*>(p));
  }
  return v;
}
// Two 128-bit streaming loads per lane for fp32 (the pre-256-bit form): the BN backward reduction, which keeps three
// input streams of four rows in flight under an 85-register cap, measured 6 % FASTER with it than with LDG.256.
template <typename T>
__device__ __forceinline__ V8<T> ld8s_split(const T* p) {
  V8<T> v;
#pragma unroll
  for (int j = 0; j < (int)sizeof(T) / 2; ++j) v.q[j] = ldg128_stream(reinterpret_cast<const uint4*>(p) + j);
  return v;
}
template <typename T>
__device__ __forceinline__ void st8(T* p, const V8<T>& v) {
  if constexpr (sizeof(T) == 4) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v.q[0].x), "r"(v.q[0].y), "r"(v.q[0].z), "r"(v.q[0].w),
                 "r"(v.q[1].x), "r"(v.q[1].y), "r"(v.q[1].z), "r"(v.q[1].w)
                 : "memory");
  } else {
    reinterpret_cast<uint4*>(p)[0] = v.q[0];
  }
}
// fp32 activations are accessed 256 bits at a time: every activation pointer of such a call must be 32-byte aligned
#define PDB_CHECK_ALIGN32(dtype, who, ...)                                                                        \
  do {                                                                                                            \
    if ((dtype) == PDB_FP32) {                                                                                    \
      const void* ptrs__[] = {__VA_ARGS__};                                                                       \
      for (const void* q__ : ptrs__)                                                                              \
        if ((reinterpret_cast<uintptr_t>(q__) & 31) != 0) {                                                        \
          ::pdb::set_error("%s: fp32 activations must be 32-byte aligned (256-bit accesses), got %p", who, q__);  \
          return 1;                                                                                               \
        }                                                                                                         \
    }                                                                                                             \
  } while (0)
template <typename T>
__device__ __forceinline__ V8<T> zero8() {
  V8<T> v;
#pragma unroll
  for (int j = 0; j < (int)sizeof(T) / 2; ++j) v.q[j] = make_uint4(0, 0, 0, 0);
  return v;
}
template <typename T>
__device__ __forceinline__ void unpack8(const V8<T>& v, float (&f)[8]) {
  if constexpr (sizeof(T) == 4) {
    const float* s = reinterpret_cast<const float*>(&v);
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = s[i];
  } else {
    unpack8<T>(v.q[0], f);
  }
}
template <typename T>
__device__ __forceinline__ V8<T> pack8v(const float (&f)[8]) {
  V8<T> v;
  if constexpr (sizeof(T) == 4) {
    float* s = reinterpret_cast<float*>(&v);
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = f[i];
  } else {
    v.q[0] = pack8<T>(f);
  }
  return v;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// dispatch on the 16-bit dtype tag of the C ABI
#define PDB_DISPATCH_16(dtype, T, ...)                         \
  do {                                                         \
    if ((dtype) == PDB_BF16) {                                 \
      using T = __nv_bfloat16;                                 \
      __VA_ARGS__;                                             \
    } else if ((dtype) == PDB_FP16) {                          \
      using T = __half;                                        \
      __VA_ARGS__;                                             \
    } else {                                                   \
      ::pdb::set_error("unsupported dtype tag %d", (int)(dtype)); \
      return 1;                                                \
    }                                                          \
  } while (0)

// dispatch on any activation dtype tag: the two 16-bit formats or fp32 storage (TF32 math in the GEMMs)
#define PDB_DISPATCH_ACT(dtype, T, ...)                        \
  do {                                                         \
    if ((dtype) == PDB_BF16) {                                 \
      using T = __nv_bfloat16;                                 \
      __VA_ARGS__;                                             \
    } else if ((dtype) == PDB_FP16) {                          \
      using T = __half;                                        \
      __VA_ARGS__;                                             \
    } else if ((dtype) == PDB_FP32) {                          \
      using T = float;                                         \
      __VA_ARGS__;                                             \
    } else {                                                   \
      ::pdb::set_error("unsupported dtype tag %d", (int)(dtype)); \
      return 1;                                                \
    }                                                          \
  } while (0)

}  // namespace pdb
