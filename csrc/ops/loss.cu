// K10: fused, loss-scaled cross-entropy forward+backward; K14: top-k hit counters.
// One warp per row (1000 classes = ~31 elements per lane), fp32 math, exp2f with pre-scaled inputs.
#include "common.cuh"

namespace pdb {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// TD: storage type of the gradient.  fp32 (the default of the Python layer) keeps (softmax - onehot)/rows exact until
// the amp loss scale has been applied: rounded to fp16 first, entries below 2^-24 * rows vanish and a scale of 65536
// itself does not fit the format.
template <typename T, typename TD>
__global__ void __launch_bounds__(256) xent_kernel(const T* __restrict__ logits, const int64_t* __restrict__ labels,
                                                   int rows, int classes, float* __restrict__ loss,
                                                   TD* __restrict__ dlogits, const float* __restrict__ grad_scale) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const T* row = logits + (int64_t)warp * classes;
  float m = -INFINITY;
  for (int c = lane; c < classes; c += 32) m = fmaxf(m, Elem<T>::to_f(row[c]));
  m = warp_max(m);
  float s = 0.f;
  for (int c = lane; c < classes; c += 32) s += exp2f((Elem<T>::to_f(row[c]) - m) * kLog2e);
  s = warp_sum(s);
  const int y = (int)labels[warp];
  const float inv_rows = 1.0f / rows;
  if (lane == 0) {
    const float xy = (y >= 0 && y < classes) ? Elem<T>::to_f(row[y]) : m;
    atomicAdd(loss, (log2f(s) * kLn2 + m - xy) * inv_rows);
  }
  if (dlogits != nullptr) {
    const float gs = (grad_scale ? *grad_scale : 1.0f) * inv_rows;
    const float inv_s = 1.0f / s;
    TD* drow = dlogits + (int64_t)warp * classes;
    for (int c = lane; c < classes; c += 32) {
      float p = exp2f((Elem<T>::to_f(row[c]) - m) * kLog2e) * inv_s;
      drow[c] = Elem<TD>::from_f((p - (c == y ? 1.f : 0.f)) * gs);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) topk_kernel(const T* __restrict__ logits, const int64_t* __restrict__ labels,
                                                   int rows, int classes, const int* __restrict__ ks, int nk,
                                                   float* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const T* row = logits + (int64_t)warp * classes;
  const int y = (int)labels[warp];
  if (y < 0 || y >= classes) return;
  const float xy = Elem<T>::to_f(row[y]);
  int rank = 0;     // number of classes that sort before the label
  for (int c = lane; c < classes; c += 32) {
    const float v = Elem<T>::to_f(row[c]);
    rank += (v > xy) || (v == xy && c < y);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) rank += __shfl_xor_sync(0xffffffffu, rank, o);
  if (lane == 0)
    for (int j = 0; j < nk; ++j)
      if (rank < ks[j]) atomicAdd(out + j, 1.0f);
}

}  // namespace pdb
using namespace pdb;

extern "C" int pdb_xent_fwd_bwd(const void* logits, int dtype, const int64_t* labels, int rows, int classes,
                                float* loss, void* dlogits, int dlogits_dtype, const float* grad_scale, void* stream) {
  PDB_CHECK_ARG(logits && labels && loss && rows > 0 && classes > 0, "pdb_xent_fwd_bwd: bad arguments");
  PDB_CHECK_ARG(!dlogits || dlogits_dtype == dtype || dlogits_dtype == PDB_FP32,
                "pdb_xent_fwd_bwd: dlogits must have the logits' dtype or be fp32");
  const int grid = (rows * 32 + 255) / 256;
  cudaStream_t st = (cudaStream_t)stream;
  if (dlogits && dlogits_dtype == PDB_FP32)
    PDB_DISPATCH_ACT(dtype, T, (xent_kernel<T, float><<<grid, 256, 0, st>>>((const T*)logits, labels, rows, classes, loss,
                                                                            (float*)dlogits, grad_scale)));
  else
    PDB_DISPATCH_ACT(dtype, T, (xent_kernel<T, T><<<grid, 256, 0, st>>>((const T*)logits, labels, rows, classes, loss,
                                                                        (T*)dlogits, grad_scale)));
  PDB_LAUNCH_CHECK();
  return 0;
}

extern "C" int pdb_topk_correct(const void* logits, int dtype, const int64_t* labels, int rows, int classes,
                                const int* ks, int nk, float* out, void* stream) {
  PDB_CHECK_ARG(logits && labels && ks && out && rows > 0 && classes > 0 && nk > 0, "pdb_topk_correct: bad arguments");
  const int grid = (rows * 32 + 255) / 256;
  PDB_DISPATCH_ACT(dtype, T, (topk_kernel<T><<<grid, 256, 0, (cudaStream_t)stream>>>((const T*)logits, labels, rows,
                                                                                     classes, ks, nk, out)));
  PDB_LAUNCH_CHECK();
  return 0;
}
