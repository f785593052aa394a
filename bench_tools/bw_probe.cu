// HBM bandwidth probe for the elementwise (BatchNorm-like) kernels: which loop shape / cache hints get
// closest to the measured copy bandwidth for a 2-read + 1-write 16-bit stream?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/bw_probe bench_tools/bw_probe.cu && build/bw_probe
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_plain(const void* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void st_plain(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }
__device__ __forceinline__ void st_cs(void* p, uint4 v) {
  asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 combine(uint4 a, uint4 b) {
  __nv_bfloat162* x = reinterpret_cast<__nv_bfloat162*>(&a);
  __nv_bfloat162* y = reinterpret_cast<__nv_bfloat162*>(&b);
#pragma unroll
  for (int i = 0; i < 4; ++i) x[i] = __hmax2(__hfma2(x[i], y[i], y[i]), __float2bfloat162_rn(0.f));
  return a;
}

// MODE: 0 = 2R+1W, 1 = copy (1R+1W), 2 = 3R reduce (no write)
template <int U, bool STRIDE, bool STREAM_LD, bool CS_ST, int MODE>
__global__ void k(const uint4* __restrict__ a, const uint4* __restrict__ b, const uint4* __restrict__ c, uint4* __restrict__ o,
                  long long n, unsigned* sink) {
  const long long tile = (long long)blockDim.x * U;
  unsigned acc = 0;
  for (long long base = (long long)blockIdx.x * tile; base < n; base += (long long)gridDim.x * tile) {
    uint4 va[U], vb[U], vc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = base + (long long)u * blockDim.x + threadIdx.x;
      if (i < n) {
        va[u] = STREAM_LD ? ld_stream(a + i) : ld_plain(a + i);
        if (MODE != 1) vb[u] = STREAM_LD ? ld_stream(b + i) : ld_plain(b + i);
        if (MODE == 2) vc[u] = STREAM_LD ? ld_stream(c + i) : ld_plain(c + i);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = base + (long long)u * blockDim.x + threadIdx.x;
      if (i < n) {
        if (MODE == 2) {
          uint4 t = combine(combine(va[u], vb[u]), vc[u]);
          acc += t.x ^ t.y ^ t.z ^ t.w;
        } else {
          uint4 r = MODE == 1 ? va[u] : combine(va[u], vb[u]);
          if (CS_ST) st_cs(o + i, r); else st_plain(o + i, r);
        }
      }
    }
    if (!STRIDE) break;
  }
  if (MODE == 2 && acc == 0x12345678u) *sink = acc;
}

template <typename F>
float timeit(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) f();
  cudaEventRecord(e0);
  for (int i = 0; i < 10; ++i) f();
  cudaEventRecord(e1);
  CK(cudaEventSynchronize(e1));
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  return ms / 10;
}

int main() {
  const long long bytes = 400ll * 56 * 56 * 256 * 2;          // one 56x56x256 activation at batch 400
  const long long n = bytes / 16;
  uint4 *a, *b, *c, *o;
  unsigned* sink;
  CK(cudaMalloc(&a, bytes));
  CK(cudaMalloc(&b, bytes));
  CK(cudaMalloc(&c, bytes));
  CK(cudaMalloc(&o, bytes));
  CK(cudaMalloc(&sink, 4));
  CK(cudaMemset(a, 0x3c, bytes));
  CK(cudaMemset(b, 0x3c, bytes));
  CK(cudaMemset(c, 0x3c, bytes));
  int sms;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  auto report = [&](const char* name, float ms, int streams) {
    printf("%-44s %8.3f ms  %7.1f GB/s\n", name, ms, streams * (double)bytes / ms / 1e6);
  };
#define RUN(NAME, U, STRIDE, SLD, CS, MODE, THREADS, GRID, STREAMS)                                          \
  report(NAME, timeit([&] { k<U, STRIDE, SLD, CS, MODE><<<GRID, THREADS>>>(a, b, c, o, n, sink); }), STREAMS)
  const int full256_4 = (int)((n + 256 * 4 - 1) / (256 * 4)), full512_4 = (int)((n + 512 * 4 - 1) / (512 * 4));
  const int full256_8 = (int)((n + 256 * 8 - 1) / (256 * 8)), full256_2 = (int)((n + 256 * 2 - 1) / (256 * 2));
  report("cudaMemcpy D2D (1R+1W)", timeit([&] { cudaMemcpyAsync(o, a, bytes, cudaMemcpyDeviceToDevice); }), 2);
  RUN("copy  stride U4 256thr grid 8/SM", 4, true, true, false, 1, 256, sms * 8, 2);
  RUN("copy  tiled  U4 256thr", 4, false, true, false, 1, 256, full256_4, 2);
  RUN("2R1W  stride U4 256thr grid 8/SM (current)", 4, true, true, false, 0, 256, sms * 8, 3);
  RUN("2R1W  stride U4 256thr grid 4/SM", 4, true, true, false, 0, 256, sms * 4, 3);
  RUN("2R1W  stride U4 256thr grid 8/SM st.cs", 4, true, true, true, 0, 256, sms * 8, 3);
  RUN("2R1W  stride U4 256thr grid 8/SM plain ld", 4, true, false, false, 0, 256, sms * 8, 3);
  RUN("2R1W  stride U8 256thr grid 4/SM", 8, true, true, false, 0, 256, sms * 4, 3);
  RUN("2R1W  stride U2 256thr grid 8/SM", 2, true, true, false, 0, 256, sms * 8, 3);
  RUN("2R1W  stride U4 512thr grid 4/SM", 4, true, true, false, 0, 512, sms * 4, 3);
  RUN("2R1W  stride U4 1024thr grid 2/SM", 4, true, true, false, 0, 1024, sms * 2, 3);
  RUN("2R1W  tiled  U4 256thr", 4, false, true, false, 0, 256, full256_4, 3);
  RUN("2R1W  tiled  U4 256thr st.cs", 4, false, true, true, 0, 256, full256_4, 3);
  RUN("2R1W  tiled  U2 256thr", 2, false, true, false, 0, 256, full256_2, 3);
  RUN("2R1W  tiled  U8 256thr", 8, false, true, false, 0, 256, full256_8, 3);
  RUN("2R1W  tiled  U4 512thr", 4, false, true, false, 0, 512, full512_4, 3);
  RUN("2R1W  tiled  U4 256thr plain ld", 4, false, false, false, 0, 256, full256_4, 3);
  RUN("3R    stride U4 256thr grid 4/SM", 4, true, true, false, 2, 256, sms * 4, 3);
  RUN("3R    tiled  U4 256thr", 4, false, true, false, 2, 256, full256_4, 3);
  return 0;
}
