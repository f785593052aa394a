// K9: fused SGD (momentum, weight decay, un-scale, device-side skip, 16-bit shadow, grad zero).
// One pass over flat arenas: reads p, g, buf; writes p, buf (+ shadow, + zeroed g).  Memory-bound:
// 5-6 fp32 streams per element, 128-bit accesses, grid sized to a multiple of the SM count.
#include "common.cuh"

namespace pdb {

struct SgdArgs {
  float lr, momentum, dampening, wd;
  int nesterov, maximize, first_step, zero_grad;
};

template <typename T, bool kShadow>
__global__ void __launch_bounds__(256) sgd_kernel(float* __restrict__ p, float* __restrict__ g,
                                                  float* __restrict__ buf, T* __restrict__ shadow, int64_t n4,
                                                  SgdArgs a, const float* __restrict__ inv_scale,
                                                  const float* __restrict__ found_inf) {
  const bool skip = found_inf != nullptr && *found_inf != 0.0f;
  float s = inv_scale != nullptr ? *inv_scale : 1.0f;
  if (a.maximize) s = -s;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 gv = reinterpret_cast<const float4*>(g)[i];
    if (a.zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (skip) continue;
    float4 pv = reinterpret_cast<const float4*>(p)[i];
    float gg[4] = {gv.x * s, gv.y * s, gv.z * s, gv.w * s};
    float pp[4] = {pv.x, pv.y, pv.z, pv.w};
    float bb[4] = {0.f, 0.f, 0.f, 0.f};
    if (buf != nullptr && !a.first_step) {
      float4 bv = reinterpret_cast<const float4*>(buf)[i];
      bb[0] = bv.x, bb[1] = bv.y, bb[2] = bv.z, bb[3] = bv.w;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float d = fmaf(a.wd, pp[k], gg[k]);
      if (buf != nullptr) {
        bb[k] = a.first_step ? d : fmaf(a.momentum, bb[k], (1.0f - a.dampening) * d);
        d = a.nesterov ? fmaf(a.momentum, bb[k], d) : bb[k];
      }
      pp[k] = fmaf(-a.lr, d, pp[k]);
    }
    reinterpret_cast<float4*>(p)[i] = make_float4(pp[0], pp[1], pp[2], pp[3]);
    if (buf != nullptr) reinterpret_cast<float4*>(buf)[i] = make_float4(bb[0], bb[1], bb[2], bb[3]);
    if (kShadow) {
      using T2 = typename Elem<T>::T2;
      T2 lo = Elem<T>::from_f2(pp[0], pp[1]), hi = Elem<T>::from_f2(pp[2], pp[3]);
      uint2 w;
      w.x = *reinterpret_cast<uint32_t*>(&lo);
      w.y = *reinterpret_cast<uint32_t*>(&hi);
      reinterpret_cast<uint2*>(shadow)[i] = w;
    }
  }
}

__global__ void __launch_bounds__(256) check_finite_kernel(const float* __restrict__ g, int64_t n4,
                                                           float* __restrict__ found_inf) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  bool bad = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 v = __ldg(reinterpret_cast<const float4*>(g) + i);
    // finite <=> exponent bits not all ones
    bad |= !(isfinite(v.x) && isfinite(v.y) && isfinite(v.z) && isfinite(v.w));
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) *found_inf = 1.0f;
}

// K13: the amp scale update (torch:amp/grad_scaler.py:484-559, ATen amp_update_scale) as one thread on the device:
// back off on a non-finite step, grow after `interval` clean steps (never to a non-finite scale), refresh 1/scale
// (computed in double like torch's unscale_), clear the flag for the next step.  No host round trip.
__global__ void scaler_update_kernel(float* scale, int* tracker, float* found_inf, float* inv_scale, float growth,
                                     float backoff, int interval) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  float s = *scale;
  int t = *tracker;
  if (*found_inf > 0.f) {
    s *= backoff;
    t = 0;
  } else if (++t >= interval) {
    const float grown = s * growth;
    if (isfinite(grown)) s = grown;
    t = 0;
  }
  *scale = s;
  *tracker = t;
  *inv_scale = (float)(1.0 / (double)s);
  *found_inf = 0.f;
}

template <typename T>
__global__ void __launch_bounds__(256) cast_kernel(const float* __restrict__ src, T* __restrict__ dst, int64_t n4) {
  using T2 = typename Elem<T>::T2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 v = __ldg(reinterpret_cast<const float4*>(src) + i);
    T2 lo = Elem<T>::from_f2(v.x, v.y), hi = Elem<T>::from_f2(v.z, v.w);
    uint2 w;
    w.x = *reinterpret_cast<uint32_t*>(&lo);
    w.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(dst)[i] = w;
  }
}

static inline int flat_grid(int64_t n4) {
  int64_t blocks = (n4 + 255) / 256;
  int64_t cap = (int64_t)num_sms() * 8;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace pdb

extern "C" int pdb_sgd_step(float* p, float* g, float* buf, void* shadow16, int shadow_dtype, int64_t n, float lr,
                            float momentum, float dampening, float weight_decay, int nesterov, int maximize,
                            int first_step, const float* inv_scale, const float* found_inf, int zero_grad,
                            void* stream) {
  using namespace pdb;
  PDB_CHECK_ARG(p && g && n > 0, "pdb_sgd_step: null arena or n <= 0");
  PDB_CHECK_ARG(n % 4 == 0, "pdb_sgd_step: n (%lld) must be a multiple of 4 (arenas are padded)", (long long)n);
  PDB_CHECK_ARG(((uintptr_t)p | (uintptr_t)g | (uintptr_t)buf | (uintptr_t)shadow16) % 16 == 0,
                "pdb_sgd_step: arenas must be 16-byte aligned");
  PDB_CHECK_ARG(momentum == 0.f || buf != nullptr, "pdb_sgd_step: momentum != 0 needs a momentum buffer");
  SgdArgs a{lr, momentum, dampening, weight_decay, nesterov, maximize, first_step, zero_grad};
  if (momentum == 0.f) buf = nullptr;
  const int64_t n4 = n / 4;
  cudaStream_t st = (cudaStream_t)stream;
  if (shadow16 == nullptr) {
    sgd_kernel<__nv_bfloat16, false><<<flat_grid(n4), 256, 0, st>>>(p, g, buf, nullptr, n4, a, inv_scale, found_inf);
  } else {
    PDB_DISPATCH_16(shadow_dtype, T,
                    (sgd_kernel<T, true><<<flat_grid(n4), 256, 0, st>>>(p, g, buf, (T*)shadow16, n4, a, inv_scale, found_inf)));
  }
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_check_finite(const float* g, int64_t n, float* found_inf, void* stream) {
  using namespace pdb;
  PDB_CHECK_ARG(g && found_inf && n > 0 && n % 4 == 0, "pdb_check_finite: bad arguments");
  check_finite_kernel<<<flat_grid(n / 4), 256, 0, (cudaStream_t)stream>>>(g, n / 4, found_inf);
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_scaler_update(float* scale, int* growth_tracker, float* found_inf, float* inv_scale, float growth,
                                 float backoff, int interval, void* stream) {
  using namespace pdb;
  PDB_CHECK_ARG(scale && growth_tracker && found_inf && inv_scale && interval > 0, "pdb_scaler_update: bad arguments");
  scaler_update_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(scale, growth_tracker, found_inf, inv_scale, growth, backoff,
                                                           interval);
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_cast_f32_to_16(const float* src, void* dst, int dtype, int64_t n, void* stream) {
  using namespace pdb;
  PDB_CHECK_ARG(src && dst && n > 0 && n % 4 == 0, "pdb_cast_f32_to_16: bad arguments");
  PDB_DISPATCH_16(dtype, T, (cast_kernel<T><<<flat_grid(n / 4), 256, 0, (cudaStream_t)stream>>>(src, (T*)dst, n / 4)));
  PDB_LAUNCH_CHECK();
  return 0;
}
