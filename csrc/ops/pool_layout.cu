// K12 pooling and the layout / elementwise helpers, NHWC 16-bit, 8 channels per thread (128-bit).
#include "common.cuh"

namespace pdb {

static int ew_grid(int64_t n, int threads = 256) {
  int64_t want = (n + threads - 1) / threads;
  int64_t cap = (int64_t)num_sms() * 8;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

// ---- MaxPool 3x3 / stride 2 / pad 1 -------------------------------------------------------------
// argmax stores the window position 0..8 (kh*3+kw) of the first maximum, like cuDNN/ATen, so the
// backward pass routes the gradient to exactly one input even when 16-bit values tie.
// One block per output row (n, ho): the index arithmetic is one 32-bit division per element instead of the chain of
// 64-bit divisions a flat index needs (those, not memory, bounded the first version: 0.37 ms for 0.9 GB).
template <typename T>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                          uint8_t* __restrict__ argmax, int H, int W, int C,
                                                          int Ho, int Wo) {
  const int cvec = C / 8;
  const int n = blockIdx.x / Ho, ho = blockIdx.x - n * Ho;
  const T* xn = x + (int64_t)n * H * W * C;
  const int64_t orow = ((int64_t)n * Ho + ho) * Wo * C;
  for (int j = threadIdx.x; j < Wo * cvec; j += blockDim.x) {
    const int wo = j / cvec, cg = j - wo * cvec;
    float best[8];
    int idx[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) best[k] = -INFINITY, idx[k] = 0;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int h = 2 * ho - 1 + kh;
      if (h < 0 || h >= H) continue;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int w = 2 * wo - 1 + kw;
        if (w < 0 || w >= W) continue;
        float f[8];
        unpack8<T>(ld8<T>(xn + ((int64_t)h * W + w) * C + cg * 8), f);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (f[k] > best[k]) best[k] = f[k], idx[k] = kh * 3 + kw;
      }
    }
    const int64_t o = orow + (int64_t)j * 8;
    st8<T>(y + o, pack8v<T>(best));
    uint2 packed;
    packed.x = idx[0] | (idx[1] << 8) | (idx[2] << 16) | (idx[3] << 24);
    packed.y = idx[4] | (idx[5] << 8) | (idx[6] << 16) | (idx[7] << 24);
    *reinterpret_cast<uint2*>(argmax + o) = packed;
  }
}

// Gather form: every input pixel collects from the (at most 2x2) windows that contain it — no atomics.
// One block per input row (n, h).
template <typename T>
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ argmax,
                                                          T* __restrict__ dx, int H, int W, int C, int Ho, int Wo) {
  const int cvec = C / 8;
  const int n = blockIdx.x / H, h = blockIdx.x - n * H;
  // windows with 2*ho-1 <= h <= 2*ho+1  <=>  ho in [ceil((h-1)/2) .. floor((h+1)/2)] = [h/2 .. (h+1)/2]
  const int ho0 = h / 2, ho1 = min((h + 1) / 2, Ho - 1);
  const int64_t obase = (int64_t)n * Ho * Wo * C;
  T* dxrow = dx + ((int64_t)n * H + h) * W * C;
  for (int j = threadIdx.x; j < W * cvec; j += blockDim.x) {
    const int w = j / cvec, cg = j - w * cvec;
    const int wo0 = w / 2, wo1 = min((w + 1) / 2, Wo - 1);
    // the (up to) four windows are loaded up front — independent loads in flight instead of a dependent loop — and
    // the argmax bytes are compared four at a time
    uint2 a[4];
    V8<T> g[4];
    uint32_t pos4[4];
    bool live[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int ho = (t >> 1) ? ho1 : ho0, wo = (t & 1) ? wo1 : wo0;
      live[t] = !((t >> 1) && ho1 == ho0) && !((t & 1) && wo1 == wo0);
      pos4[t] = (uint32_t)((h - (2 * ho - 1)) * 3 + (w - (2 * wo - 1))) * 0x01010101u;
      if (live[t]) {
        const int64_t o = obase + ((int64_t)ho * Wo + wo) * C + cg * 8;
        a[t] = __ldg(reinterpret_cast<const uint2*>(argmax + o));
        g[t] = ld8<T>(dy + o);
      }
    }
    float acc[8] = {0};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      if (!live[t]) continue;
      const uint32_t m0 = __vcmpeq4(a[t].x, pos4[t]), m1 = __vcmpeq4(a[t].y, pos4[t]);   // 0xff per matching byte
      float f[8];
      unpack8<T>(g[t], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += ((k < 4 ? m0 >> (8 * k) : m1 >> (8 * (k - 4))) & 1u) ? f[k] : 0.f;
    }
    st8<T>(dxrow + (int64_t)j * 8, pack8v<T>(acc));
  }
}

// ---- global average pool [N, HW, C] -> [N, C] ---------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int N, int HW,
                                                          int C) {
  const int cvec = C / 8;
  const int64_t total = (int64_t)N * cvec, stride = (int64_t)gridDim.x * blockDim.x;
  const float inv = 1.0f / HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cg = (int)(i % cvec);
    const int n = (int)(i / cvec);
    float acc[8] = {0};
    for (int p = 0; p < HW; ++p) {
      float f[8];
      unpack8<T>(ld8<T>(x + ((int64_t)n * HW + p) * C + cg * 8), f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] *= inv;
    st8<T>(y + i * 8, pack8v<T>(acc));
  }
}

template <typename T>
__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int N, int HW,
                                                          int C) {
  const int cvec = C / 8;
  const int64_t total = (int64_t)N * HW * cvec, stride = (int64_t)gridDim.x * blockDim.x;
  const float inv = 1.0f / HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cg = (int)(i % cvec);
    const int n = (int)(i / ((int64_t)HW * cvec));
    float f[8];
    unpack8<T>(ld8<T>(dy + ((int64_t)n * C + cg * 8)), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] *= inv;
    st8<T>(dx + i * 8, pack8v<T>(f));
  }
}

// ---- NCHW fp32 -> NHWC 16-bit (C small, e.g. 3): one thread per pixel ---------------------------
template <typename T>
__global__ void __launch_bounds__(256) nchw_to_nhwc_kernel(const float* __restrict__ x, T* __restrict__ y, int N, int C,
                                                           int HW) {
  const int64_t total = (int64_t)N * HW, stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t n = i / HW, p = i % HW;
    for (int c = 0; c < C; ++c) y[i * C + c] = Elem<T>::from_f(__ldg(x + (n * C + c) * HW + p));
  }
}

// ---- stem im2col: [N,3,H,W] fp32 -> [N*Ho*Wo, Kpad] 16-bit, column = (kh*7 + kw)*3 + c -----------
// One block per output row (n, ho): the 7 input rows x 3 channels it needs are staged in shared memory with
// coalesced loads (zero halo included), then every thread assembles 128-bit output vectors from smem through a
// per-column offset table — no per-element div/mod, no scattered global loads, fully coalesced 34 KB of writes
// per block.  (The first version did the gather straight from global memory: 1.76 ms; this one is bandwidth
// bound on the 1.5 GB it writes.)
constexpr int kStemTaps = 7, kStemK = 147, kStemHalo = 3;
template <typename T>
__global__ void __launch_bounds__(256) stem_im2col_kernel(const float* __restrict__ x, T* __restrict__ col, int N,
                                                          int H, int W, int Ho, int Wo, int Kpad) {
  extern __shared__ float srows[];                 // [7*3][pitch] input rows with a zero halo of 3 on both sides
  const int pitch = W + 2 * kStemHalo + 2;
  const int n = blockIdx.x / Ho, ho = blockIdx.x - n * Ho;
  const int64_t HW = (int64_t)H * W;
  for (int i = threadIdx.x; i < kStemTaps * 3 * pitch; i += blockDim.x) {
    const int rc = i / pitch, wp = i - rc * pitch;           // rc = kh*3 + c
    const int kh = rc / 3, c = rc - kh * 3;
    const int h = 2 * ho - kStemHalo + kh, w = wp - kStemHalo;
    float v = 0.f;
    if (h >= 0 && h < H && w >= 0 && w < W) v = __ldg(x + ((int64_t)n * 3 + c) * HW + (int64_t)h * W + w);
    srows[i] = v;
  }
  // a thread keeps ONE 8-column group of the output row for the whole block: its eight smem offsets live in
  // registers, and it walks the pixels with a fixed stride (the first version looked the offsets up in a shared
  // table for every vector: twice the shared-memory loads, 2-byte lanes that conflict)
  const int kvec = Kpad / 8;
  const int wo_step = blockDim.x / kvec;
  const int kv = threadIdx.x % kvec, wo_first = threadIdx.x / kvec;
  int off[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int kk = kv * 8 + k;
    const int c = kk % 3, tap = kk / 3, kw = tap % kStemTaps, kh = tap / kStemTaps;
    off[k] = kk < kStemK ? (kh * 3 + c) * pitch + kw : -1;
  }
  __syncthreads();
  if (wo_first >= wo_step) return;
  T* out = col + ((int64_t)n * Ho + ho) * Wo * Kpad;
  for (int wo = wo_first; wo < Wo; wo += wo_step) {
    const float* base = srows + 2 * wo;                      // window of pixel wo starts at smem column 2*wo
    float f[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = off[k] >= 0 ? base[off[k]] : 0.f;
    st8<T>(out + ((int64_t)wo * kvec + kv) * 8, pack8v<T>(f));
  }
}

template <typename T>
__global__ void __launch_bounds__(256) add16_kernel(const T* __restrict__ a, const T* __restrict__ b,
                                                    T* __restrict__ o, int64_t nvec) {
  constexpr int U = sizeof(T) == 4 ? 4 : 8;          // one tile per block (see bn_fast.cuh for the measurement behind this shape)
  const int64_t base = (int64_t)blockIdx.x * (256 * U) + threadIdx.x;
  V8<T> va[U], vb[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int64_t i = base + (int64_t)u * 256;
    if (i < nvec) va[u] = ld8s<T>(a + i * 8), vb[u] = ld8s<T>(b + i * 8);
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int64_t i = base + (int64_t)u * 256;
    if (i >= nvec) continue;
    float fa[8], fb[8];
    unpack8<T>(va[u], fa);
    unpack8<T>(vb[u], fb);
#pragma unroll
    for (int k = 0; k < 8; ++k) fa[k] += fb[k];
    st8<T>(o + i * 8, pack8v<T>(fa));
  }
}

// dst[n, 2ho, 2wo] = src[n, ho, wo]; everything else zero.  One block per destination row (n, h).
template <typename T>
__global__ void __launch_bounds__(256) upsample2_kernel(const T* __restrict__ src, T* __restrict__ dst, int Ho, int Wo,
                                                        int C) {
  const int cvec = C / 8, H = 2 * Ho, W = 2 * Wo;
  const int n = blockIdx.x / H, h = blockIdx.x - n * H;
  T* drow = dst + ((int64_t)n * H + h) * W * C;
  const T* srow = src + ((int64_t)n * Ho + h / 2) * Wo * C;
  const bool live = (h & 1) == 0;
  for (int j = threadIdx.x; j < W * cvec; j += blockDim.x) {
    const int w = j / cvec, cg = j - w * cvec;
    V8<T> v = zero8<T>();
    if (live && (w & 1) == 0) v = ld8s<T>(srow + ((int64_t)(w / 2) * cvec + cg) * 8);
    st8<T>(drow + (int64_t)j * 8, v);
  }
}

// o = a, plus b at the even (h, w) positions (b has half the resolution).  One block per row (n, h).
template <typename T>
__global__ void __launch_bounds__(256) add_strided2_kernel(const T* __restrict__ a, const T* __restrict__ b,
                                                           T* __restrict__ o, int H, int W, int C) {
  const int cvec = C / 8, Ho = H / 2, Wo = W / 2;
  const int n = blockIdx.x / H, h = blockIdx.x - n * H;
  const int64_t row = ((int64_t)n * H + h) * W * C;
  const T* brow = b + ((int64_t)n * Ho + h / 2) * Wo * C;
  const bool live = (h & 1) == 0;
  for (int j = threadIdx.x; j < W * cvec; j += blockDim.x) {
    const int w = j / cvec, cg = j - w * cvec;
    V8<T> va = ld8s<T>(a + row + (int64_t)j * 8);
    if (live && (w & 1) == 0) {
      float fa[8], fb[8];
      unpack8<T>(va, fa);
      unpack8<T>(ld8s<T>(brow + ((int64_t)(w / 2) * cvec + cg) * 8), fb);
#pragma unroll
      for (int k = 0; k < 8; ++k) fa[k] += fb[k];
      va = pack8v<T>(fa);
    }
    st8<T>(o + row + (int64_t)j * 8, va);
  }
}

// ---- space-to-depth stem helpers (see pdb_stem_fprop) -------------------------------------------------------------
// xs[n, p, q, (dy*2 + dx)*3 + c] = x[n, c, 2p + dy, 2q + dx]; channels 12..15 are zero.  One block per output row.
template <typename T>
__global__ void __launch_bounds__(128) stem_s2d_kernel(const float* __restrict__ x, T* __restrict__ xs, int H, int W) {
  const int Hs = H / 2, Ws = W / 2;
  const int n = blockIdx.x / Hs, p = blockIdx.x - n * Hs;
  const int64_t HW = (int64_t)H * W;
  const float* xn = x + (int64_t)n * 3 * HW + (int64_t)(2 * p) * W;
  T* out = xs + ((int64_t)n * Hs + p) * Ws * 16;
  for (int q = threadIdx.x; q < Ws; q += blockDim.x) {
    float f[16];
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float2 v = __ldg(reinterpret_cast<const float2*>(xn + c * HW + dy * W + 2 * q));
        f[(dy * 2 + 0) * 3 + c] = v.x, f[(dy * 2 + 1) * 3 + c] = v.y;
      }
    f[12] = f[13] = f[14] = f[15] = 0.f;
    float lo[8], hi[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) lo[k] = f[k], hi[k] = f[8 + k];
    st8<T>(out + q * 16, pack8v<T>(lo));
    st8<T>(out + q * 16 + 8, pack8v<T>(hi));
  }
}

// ws[k, (r*4 + s)*16 + (dy*2 + dx)*3 + c] = w[k, kh, kw, c] with kh = 2r + dy - 1, kw = 2s + dx - 1 (zero outside 0..6)
template <typename T>
__global__ void stem_weight_s2d_kernel(const T* __restrict__ w, T* __restrict__ ws, int K) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= K * 256) return;
  const int k = i >> 8, col = i & 255, tap = col >> 4, ch = col & 15;
  const int r = tap >> 2, s = tap & 3, d = ch / 3, c = ch - d * 3;
  const int kh = 2 * r + (d >> 1) - 1, kw = 2 * s + (d & 1) - 1;
  T v = Elem<T>::from_f(0.f);
  if (ch < 12 && kh >= 0 && kh < 7 && kw >= 0 && kw < 7) v = w[k * 147 + (kh * 7 + kw) * 3 + c];
  ws[i] = v;
}

// dw[k, kh, kw, c] (+)= dws[k, column of (kh, kw, c)]
__global__ void stem_dw_unpack_kernel(const float* __restrict__ dws, float* __restrict__ dw, int K, int accumulate) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= K * 147) return;
  const int k = i / 147, j = i - k * 147, tapc = j / 3, c = j - tapc * 3, kh = tapc / 7, kw = tapc - kh * 7;
  const int r = (kh + 1) >> 1, dy = (kh + 1) & 1, s = (kw + 1) >> 1, dx = (kw + 1) & 1;
  const float v = dws[k * 256 + (r * 4 + s) * 16 + (dy * 2 + dx) * 3 + c];
  dw[i] = accumulate ? dw[i] + v : v;
}

}  // namespace pdb
using namespace pdb;

#define ST ((cudaStream_t)stream)

extern "C" int pdb_maxpool3x3s2_fwd(const void* x, void* y, uint8_t* argmax, int dtype, int N, int H, int W, int C,
                                    void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_maxpool3x3s2_fwd", x, y);
  PDB_CHECK_ARG(x && y && argmax && N > 0 && H > 0 && W > 0 && C % 8 == 0, "pdb_maxpool3x3s2_fwd: bad arguments");
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  PDB_CHECK_ARG((int64_t)N * Ho < (1ll << 31), "pdb_maxpool3x3s2_fwd: too many rows");
  PDB_DISPATCH_ACT(dtype, T, (maxpool_fwd_kernel<T><<<N * Ho, 256, 0, ST>>>((const T*)x, (T*)y, argmax, H, W, C, Ho, Wo)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_maxpool3x3s2_bwd(const void* dy, const uint8_t* argmax, void* dx, int dtype, int N, int H, int W,
                                    int C, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_maxpool3x3s2_bwd", dy, dx);
  PDB_CHECK_ARG(dy && dx && argmax && N > 0 && H > 0 && W > 0 && C % 8 == 0, "pdb_maxpool3x3s2_bwd: bad arguments");
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  PDB_CHECK_ARG((int64_t)N * H < (1ll << 31), "pdb_maxpool3x3s2_bwd: too many rows");
  PDB_DISPATCH_ACT(dtype, T, (maxpool_bwd_kernel<T><<<N * H, 256, 0, ST>>>((const T*)dy, argmax, (T*)dx, H, W, C, Ho, Wo)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_avgpool_fwd(const void* x, void* y, int dtype, int N, int HW, int C, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_avgpool_fwd", x, y);
  PDB_CHECK_ARG(x && y && N > 0 && HW > 0 && C % 8 == 0, "pdb_avgpool_fwd: bad arguments");
  PDB_DISPATCH_ACT(dtype, T, (avgpool_fwd_kernel<T><<<ew_grid((int64_t)N * C / 8, 128), 128, 0, ST>>>((const T*)x, (T*)y,
                                                                                                      N, HW, C)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_avgpool_bwd(const void* dy, void* dx, int dtype, int N, int HW, int C, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_avgpool_bwd", dy, dx);
  PDB_CHECK_ARG(dy && dx && N > 0 && HW > 0 && C % 8 == 0, "pdb_avgpool_bwd: bad arguments");
  PDB_DISPATCH_ACT(dtype, T, (avgpool_bwd_kernel<T><<<ew_grid((int64_t)N * HW * C / 8), 256, 0, ST>>>((const T*)dy,
                                                                                                      (T*)dx, N, HW, C)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_nchw_f32_to_nhwc16(const float* x, void* y, int dtype, int N, int C, int H, int W, void* stream) {
  PDB_CHECK_ARG(x && y && N > 0 && C > 0 && H > 0 && W > 0, "pdb_nchw_f32_to_nhwc16: bad arguments");
  PDB_DISPATCH_ACT(dtype, T, (nchw_to_nhwc_kernel<T><<<ew_grid((int64_t)N * H * W), 256, 0, ST>>>(x, (T*)y, N, C, H * W)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_stem_im2col(const float* x, void* col, int dtype, int N, int H, int W, int Kpad, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_stem_im2col", col);
  PDB_CHECK_ARG(x && col && N > 0 && H > 0 && W > 0 && Kpad >= 147 && Kpad % 8 == 0, "pdb_stem_im2col: bad arguments");
  PDB_CHECK_ARG(Kpad <= 160, "pdb_stem_im2col: Kpad must be <= 160");
  const int Ho = (H + 6 - 7) / 2 + 1, Wo = (W + 6 - 7) / 2 + 1;
  const size_t smem = (size_t)7 * 3 * (W + 8) * sizeof(float);
  PDB_CHECK_ARG(smem <= 48 * 1024, "pdb_stem_im2col: image width %d too large for the row staging buffer", W);
  PDB_DISPATCH_ACT(dtype, T, (stem_im2col_kernel<T><<<N * Ho, 256, smem, ST>>>(x, (T*)col, N, H, W, Ho, Wo, Kpad)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_add16(const void* a, const void* b, void* out, int dtype, int64_t n, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_add16", a, b, out);
  PDB_CHECK_ARG(a && b && out && n > 0 && n % 8 == 0, "pdb_add16: bad arguments");
  PDB_DISPATCH_ACT(dtype, T, (add16_kernel<T><<<(int)((n / 8 + (dtype == PDB_FP32 ? 1023 : 2047)) / (dtype == PDB_FP32 ? 1024 : 2048)), 256, 0, ST>>>((const T*)a, (const T*)b, (T*)out, n / 8)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_upsample2_zero(const void* src, void* dst, int dtype, int N, int Ho, int Wo, int C, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_upsample2_zero", src, dst);
  PDB_CHECK_ARG(src && dst && N > 0 && Ho > 0 && Wo > 0 && C % 8 == 0, "pdb_upsample2_zero: bad arguments");
  PDB_CHECK_ARG((int64_t)N * Ho * 2 < (1ll << 31), "pdb_upsample2_zero: too many rows");
  PDB_DISPATCH_ACT(dtype, T, (upsample2_kernel<T><<<N * Ho * 2, 256, 0, ST>>>((const T*)src, (T*)dst, Ho, Wo, C)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_add_strided2(const void* a, const void* b, void* out, int dtype, int N, int H, int W, int C,
                                void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_add_strided2", a, b, out);
  PDB_CHECK_ARG(a && b && out && N > 0 && H % 2 == 0 && W % 2 == 0 && C % 8 == 0, "pdb_add_strided2: bad arguments");
  PDB_CHECK_ARG((int64_t)N * H < (1ll << 31), "pdb_add_strided2: too many rows");
  PDB_DISPATCH_ACT(dtype, T, (add_strided2_kernel<T><<<N * H, 256, 0, ST>>>((const T*)a, (const T*)b, (T*)out, H, W, C)));
  PDB_LAUNCH_CHECK();
  return 0;
}

// ---- weight layout helpers --------------------------------------------------------------------
namespace pdb {
template <typename T>
__global__ void __launch_bounds__(256) weight_dgrad_kernel(const T* __restrict__ w, T* __restrict__ wt, int K, int R,
                                                           int S, int C) {
  // one thread per destination element; destination [C][R][S][K] is written coalesced along K
  const int64_t total = (int64_t)C * R * S * K, stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int k = (int)(i % K);
    int64_t t = i / K;
    const int s = (int)(t % S);
    t /= S;
    const int r = (int)(t % R);
    const int c = (int)(t / R);
    wt[i] = w[(((int64_t)k * R + (R - 1 - r)) * S + (S - 1 - s)) * C + c];
  }
}

template <typename T>
__global__ void __launch_bounds__(256) pad_rows_kernel(const T* __restrict__ src, T* __restrict__ dst, int rows,
                                                       int cs, int cd) {
  const int64_t total = (int64_t)rows * cd, stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c = (int)(i % cd);
    const int64_t r = i / cd;
    dst[i] = c < cs ? src[r * cs + c] : Elem<T>::from_f(0.f);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ a, float* __restrict__ out, int rows,
                                                     int cols) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float acc = 0.f;
  for (int r = 0; r < rows; ++r) acc += Elem<T>::to_f(a[(int64_t)r * cols + c]);
  out[c] += acc;
}
}  // namespace pdb

extern "C" int pdb_weight_dgrad_layout(const void* w, void* wt, int dtype, int K, int R, int S, int C, void* stream) {
  PDB_CHECK_ARG(w && wt && K > 0 && R > 0 && S > 0 && C > 0, "pdb_weight_dgrad_layout: bad arguments");
  PDB_DISPATCH_ACT(dtype, T, (weight_dgrad_kernel<T><<<ew_grid((int64_t)K * R * S * C), 256, 0, ST>>>((const T*)w, (T*)wt,
                                                                                                      K, R, S, C)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_pad_rows16(const void* src, void* dst, int dtype, int rows, int cols_src, int cols_dst,
                              void* stream) {
  PDB_CHECK_ARG(src && dst && rows > 0 && cols_src > 0 && cols_dst >= cols_src, "pdb_pad_rows16: bad arguments");
  PDB_DISPATCH_ACT(dtype, T, (pad_rows_kernel<T><<<ew_grid((int64_t)rows * cols_dst), 256, 0, ST>>>(
                                (const T*)src, (T*)dst, rows, cols_src, cols_dst)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_colsum16(const void* a, float* out, int dtype, int rows, int cols, void* stream) {
  PDB_CHECK_ARG(a && out && rows > 0 && cols > 0, "pdb_colsum16: bad arguments");
  PDB_DISPATCH_ACT(dtype, T, (colsum_kernel<T><<<(cols + 255) / 256, 256, 0, ST>>>((const T*)a, out, rows, cols)));
  PDB_LAUNCH_CHECK();
  return 0;
}

// Batched dgrad weight layout: one launch re-arranges every convolution of the network.
// table[i] = {w_off, wt_off, K, R, S, C} (element offsets into w_base / wt_base), in device memory.
namespace pdb {
// wt[c][r][s][k] = w[k][R-1-r][S-1-s][c]: for every tap a K x C -> C x K transpose.  32 x 32 tiles through shared
// memory so that both the reads (along c) and the writes (along k) are contiguous; the first version gathered
// straight from global memory with a stride of R*S*C elements between neighbouring threads (a 32-byte sector per
// 2-byte element: 0.2 ms per step for 51 MB).
template <typename T>
__global__ void __launch_bounds__(256) weight_dgrad_batched_kernel(const T* __restrict__ w_base, T* __restrict__ wt_base,
                                                                   const int64_t* __restrict__ table) {
  __shared__ T tile[32][33];
  const int64_t* e = table + (int64_t)blockIdx.y * 6;
  const T* w = w_base + e[0];
  T* wt = wt_base + e[1];
  const int K = (int)e[2], R = (int)e[3], S = (int)e[4], C = (int)e[5];
  const int kt = (K + 31) / 32, ct = (C + 31) / 32, taps = R * S;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;           // 32 x 8 threads
  for (int t = blockIdx.x; t < taps * kt * ct; t += gridDim.x) {
    const int tap = t / (kt * ct), rem = t - tap * (kt * ct);
    const int k0 = (rem / ct) * 32, c0 = (rem % ct) * 32;
    const int r = tap / S, sx = tap - r * S;
    const int src_tap = (R - 1 - r) * S + (S - 1 - sx);
#pragma unroll
    for (int j = ty; j < 32; j += 8) {                               // rows k0 + j of the source, contiguous along c
      const int k = k0 + j, c = c0 + tx;
      if (k < K && c < C) tile[j][tx] = w[((int64_t)k * taps + src_tap) * C + c];
    }
    __syncthreads();
#pragma unroll
    for (int j = ty; j < 32; j += 8) {                               // rows c0 + j of the destination, contiguous along k
      const int c = c0 + j, k = k0 + tx;
      if (c < C && k < K) wt[((int64_t)c * taps + tap) * K + k] = tile[tx][j];
    }
    __syncthreads();
  }
}
}  // namespace pdb

extern "C" int pdb_weight_dgrad_layout_batched(const void* w_base, void* wt_base, const int64_t* table_dev, int n,
                                               int dtype, void* stream) {
  PDB_CHECK_ARG(w_base && wt_base && table_dev && n > 0 && n <= 65535, "pdb_weight_dgrad_layout_batched: bad arguments");
  dim3 grid(512, n);          // the largest layers have ~2,300 tiles: enough blocks that none walks more than a few
  PDB_DISPATCH_ACT(dtype, T, (weight_dgrad_batched_kernel<T><<<grid, 256, 0, ST>>>((const T*)w_base, (T*)wt_base, table_dev)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_stem_s2d(const float* x, void* xs, int dtype, int N, int H, int W, void* stream) {
  PDB_CHECK_ALIGN32(dtype, "pdb_stem_s2d", xs);
  PDB_CHECK_ARG(x && xs && N > 0 && H > 0 && W > 0 && H % 2 == 0 && W % 2 == 0, "pdb_stem_s2d: H and W must be even");
  PDB_CHECK_ARG((reinterpret_cast<uintptr_t>(x) & 7) == 0, "pdb_stem_s2d: input must be 8-byte aligned");
  PDB_DISPATCH_ACT(dtype, T, (stem_s2d_kernel<T><<<N * (H / 2), 128, 0, ST>>>(x, (T*)xs, H, W)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_stem_weight_s2d(const void* w, void* ws, int dtype, int K, void* stream) {
  PDB_CHECK_ARG(w && ws && K > 0, "pdb_stem_weight_s2d: bad arguments");
  PDB_DISPATCH_ACT(dtype, T, (stem_weight_s2d_kernel<T><<<(K * 256 + 255) / 256, 256, 0, ST>>>((const T*)w, (T*)ws, K)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_stem_dw_unpack(const float* dws, float* dw, int K, int accumulate, void* stream) {
  PDB_CHECK_ARG(dws && dw && K > 0, "pdb_stem_dw_unpack: bad arguments");
  stem_dw_unpack_kernel<<<(K * 147 + 255) / 256, 256, 0, ST>>>(dws, dw, K, accumulate);
  PDB_LAUNCH_CHECK();
  return 0;
}
