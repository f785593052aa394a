// K7 / K11: convolutions and the fc layer as implicit GEMMs on the 5th-generation tensor cores.
//
//   fprop / dgrad :  D[M, N] = A[M, K] * B[N, K]^T      (pdb_conv_fprop)
//       A = NHWC activation.  1x1/stride-1 (and pre-gathered) inputs are a plain [M, C] matrix loaded
//       with tiled TMA; everything else uses TMA *im2col* mode, which walks 128 consecutive output
//       pixels (wrapping rows and images), applies the filter-tap offset and zero-fills the halo.
//       B = weights [Cout, kh, kw, Cin] (K-major).  dgrad is the same kernel on dY with the flipped /
//       transposed weight copy.
//   wgrad         :  dW[Cout, K] += dY[M, Cout]^T * im2col(X)[M, K]   (pdb_conv_wgrad)
//       the reduction runs over pixels, so both operands are MN-major in shared memory; the pixel range
//       is split over CTAs and partial tiles are added to the fp32 gradient with vector reductions.
//
// Structure (both kernels): persistent CTAs, one per SM; warp 0 = TMA producer (one elected thread; its address
// arithmetic is division-free), warp 1 = MMA issuer (one elected thread issues tcgen05.mma), warps 2-9 =
// epilogue (two warps per TMEM lane quarter).  kStages-deep smem ring guarded by full/empty mbarriers; the fp32
// accumulator lives in TMEM and is double-buffered so the epilogue of tile i overlaps the main loop of tile
// i+1.  The fprop epilogue converts to 16-bit, stages at most 128 columns at a time in 128B-swizzled smem,
// writes them with TMA (plain store, or reduction store when the tile is added to an existing gradient), and
// (optionally) accumulates the per-channel sum and sum-of-squares that the following BatchNorm needs — the
// statistics pass of the delegate's BN kernel (torch:nn/modules/batchnorm.py:144-210 over cuDNN) disappears.
//
// Delegates being replaced: cuDNN 9.22 convolutions and cuBLAS(Lt) linear.
#include <cstdlib>

#include "common.cuh"
#include "tc05.cuh"
#include "comm/peer_sync.cuh"

namespace pdb {
using namespace tc05;

constexpr int BLOCK_M = 128;
// Operand formats: kFmt is the dtype tag of the C ABI (0 = bf16, 1 = fp16 -> kind::f16; 2 = fp32 storage ->
// kind::tf32).  Everything in the pipelines is sized in BYTES — a k-block is one 128-byte swizzle row per tile row, an
// MMA consumes 32 bytes of K — so the 32-bit format changes element counts only: 32 instead of 64 elements per
// k-block, K = 8 instead of 16 per MMA, and twice as many k-blocks for the same reduction length.
constexpr int kFmtTf32 = 2;
template <int kFmt>
struct Fmt {
  static constexpr int ES = kFmt == kFmtTf32 ? 4 : 2;        // bytes per element (operands and outputs)
  static constexpr int BLOCK_K = 128 / ES;                   // elements per 128-byte swizzle row: 64 / 32
  static constexpr int UMMA_K = 32 / ES;                     // 16 / 8
};
constexpr int kNumThreads = 320;            // 10 warps: producer, mma, 8 x epilogue
constexpr int kEpiThreads = 256;            // two warps per TMEM lane quarter, splitting the columns:
                                            // with K <= 512 the epilogue, not the MMA, sets the tile rate
constexpr uint32_t A_STAGE_BYTES = BLOCK_M * 128;           // 128 rows x one 128-byte swizzle row = 16 KB

// Division by a runtime constant without the ~100-cycle integer divide: the TMA producer is ONE thread and
// its address arithmetic sits on the critical path of every pipeline stage.  Valid for 0 <= n < 2^31.
struct FastDiv {
  uint32_t d, mul, shr;
  __host__ void init(uint32_t div) {
    d = div;
    if (div <= 1) {
      mul = 0, shr = 0;
      return;
    }
    uint32_t l = 0;
    while ((1u << l) < div) ++l;                 // ceil(log2(div))
    const uint32_t p = 31 + l;
    mul = (uint32_t)((((uint64_t)1 << p) + div - 1) / div);
    shr = p - 32;
  }
  __device__ __forceinline__ uint32_t div(uint32_t n) const { return d <= 1 ? n : (__umulhi(n, mul) >> shr); }
};

constexpr int kStatReplicas = 16;

// Order in which a persistent CTA (or CTA pair) walks the output tiles.  Tiles are numbered n-fastest
// (tile = m_unit * num_n_blocks + n_blk) and unit u takes u, u + G, u + 2G, ...: at any moment the grid works on
// G / num_n_blocks activation tiles, each read by num_n_blocks CTAs at about the same time, so the activation
// matrix comes from DRAM once.  The column block of unit u's k-th tile is (u + k G) mod num_n_blocks, which cycles
// with period P = num_n_blocks / gcd(G, num_n_blocks); the walk visits k = 0, P, 2P, ... then k = 1, P+1, ... so
// a unit changes column block at most P times and the per-channel statistics stay in registers in between.
struct TileWalk {
  int unit, num_units, num_tiles, period, phase, k;
  __device__ __forceinline__ TileWalk(int u, int g, int t, int p) : unit(u), num_units(g), num_tiles(t), period(p), phase(0), k(0) {}
  __device__ __forceinline__ bool next(int& tile) {
    while (phase < period) {
      tile = unit + k * num_units;
      if (tile < num_tiles) {
        k += period;
        return true;
      }
      k = ++phase;
    }
    return false;
  }
};

struct FpropParams {
  int M, N;
  int num_m_blocks, num_n_blocks;
  int tile_period;     // see TileWalk
  FastDiv div_nb;      // tile -> (m unit, column block)
  int taps, cin_blocks, S;
  int P, Q, stride, pad;
  int a_im2col;
  int accumulate;      // epilogue adds the tile into y (TMA reduction store) instead of overwriting it
  FastDiv div_pq, div_q;
  float* stats;        // [2N]: sum | sumsq, accumulated with atomics; may be null
  float* stats_rep;    // optional kStatReplicas x [2N] spread accumulators (+ ticket word after them): CTA b adds
                       // to copy b % R and the last CTA folds the copies into `stats` — hundreds of CTAs adding to
                       // the same 2N addresses serialise in L2 (~100 ns each, 20-30 us per launch otherwise)
  const float* bias;   // [N]; may be null
  SyncPeers sync;      // world > 1: SyncBatchNorm — the CTA that folds the replicas also exchanges the [2N] sums with
                       // the peer GPUs (one meeting over the signal pad), so `stats` comes back global
};

// kPair: the CTA is one half of a cta_group::2 pair working on a 256-row tile; it stages only HALF of the B tile.
// ES: bytes per output element.  The epilogue stages the tile in boxes of 128 rows x 128 B (64 16-bit or 32 fp32
// columns); a pass is at most 32 KB of them (128 16-bit or 64 fp32 columns), so that a wide tile does not spend
// pipeline stages on staging.
// kTwoOut: TWO staging buffers used alternately, so that the TMA store of pass i reads its buffer while pass i+1 is
// being converted into the other one (one buffer: every pass first waits until the previous store has drained it).
// Paid for with pipeline stages, so it is the variant for the short-K layers (<= 4 k-blocks per tile), where the
// epilogue, not the main loop, sets the tile rate and two or three stages in flight are all the main loop can use.
template <int BLOCK_N, bool kPair = false, int ES = 2, bool kTwoOut = false>
struct FpropSmem {
  static constexpr uint32_t B_STAGE_BYTES = (kPair ? BLOCK_N / 2 : BLOCK_N) * 128;
  static constexpr uint32_t STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int BOX_N = 128 / ES;                               // columns per staging box: 64 / 32
  static constexpr int PASS_MAX = 2 * BOX_N;                           // two boxes per pass
  static constexpr int PASS_N = BLOCK_N < PASS_MAX ? BLOCK_N : PASS_MAX;   // columns staged per epilogue pass
  static constexpr uint32_t OUT_BYTES = (PASS_N / BOX_N) * 16384;      // 128 rows x 128 B per box
  static constexpr int OUT_BUFS = kTwoOut ? 2 : 1;
  static constexpr int kNominal = kPair ? (BLOCK_N == 256 ? 6 : 8) : (BLOCK_N == 256 ? 4 : (BLOCK_N == 128 ? 6 : 8));
  static constexpr int kFit = (int)((232448u - 1280u - OUT_BUFS * OUT_BYTES) / STAGE_BYTES);     // 227 KB per CTA
  static constexpr int kStages = kNominal < kFit ? kNominal : kFit;
  static constexpr uint32_t BAR_OFFSET = kStages * STAGE_BYTES + OUT_BUFS * OUT_BYTES;
  static constexpr uint32_t TOTAL = BAR_OFFSET + 256 + 1024;           // barriers + alignment slack
};

template <int kFmt>
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  if (kFmt == 0) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  }
}
template <int kFmt>
__device__ __forceinline__ float2 unpack2(uint32_t w) {
  if (kFmt == 0) return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&w));
  return __half22float2(*reinterpret_cast<__half2*>(&w));
}

// kStem: the 7x7/2 stem as a 4x4/1 convolution over the space-to-depth input (16 channels per pixel, see
// pdb_stem_fprop): a pipeline stage holds one row of four filter taps, each tap a [128 pixels x 32 B] block (A) and a
// [64 couts x 32 B] block (B) in 32-byte-swizzled K-major layout; one MMA (K = 16) consumes one tap.
template <int BLOCK_N, int kFmt, bool kPair, bool kStem = false, bool kTwoOut = false>
__global__ void __launch_bounds__(kNumThreads, 1)
conv_fprop_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmD, const FpropParams p) {
  using F = Fmt<kFmt>;
  constexpr bool kTf32 = kFmt == kFmtTf32;
  constexpr int BLOCK_K = F::BLOCK_K, UMMA_K = F::UMMA_K;
  static_assert(!(kStem && kTf32), "the space-to-depth stem kernels are 16-bit only");
  static_assert(!(kStem && kTwoOut), "the stem kernel keeps one staging buffer");
  using L = FpropSmem<BLOCK_N, kPair, F::ES, kTwoOut>;
  // pair mode: rank 0 is the leader (issues the MMAs, owns the full / tmem-empty barriers that are waited on)
  const uint32_t rank = kPair ? cluster_ctarank() : 0u;
  const int unit = kPair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;        // tile-walking unit: CTA or CTA pair
  const int num_units = kPair ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int m_units = kPair ? (p.num_m_blocks + 1) / 2 : p.num_m_blocks;    // M blocks per unit column
  constexpr int kStages = L::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* out_stage = smem + kStages * L::STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + L::BAR_OFFSET);
  uint64_t* empty = full + kStages;
  uint64_t* tfull = empty + kStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    prefetch_tmap(&tmD);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(full + i, kPair ? 2 : 1);        // pair: leader's expect_tx arrive + the peer's remote arrive
      mbar_init(empty + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull + i, 1);
      mbar_init(tempty + i, (kPair ? 2 : 1) * (kEpiThreads / 32));   // one arrive per epilogue warp (of both CTAs)
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    if (kPair) tmem_alloc_2sm<2 * BLOCK_N>(tmem_slot);
    else tmem_alloc<2 * BLOCK_N>(tmem_slot);
  }
  fence_before_thread_sync();
  if (kPair) cluster_sync_all();                 // the peer's barriers must exist before anything arrives on them
  else __syncthreads();
  fence_after_thread_sync();
  const uint32_t tmem_base = *tmem_slot;

  const int num_tiles = m_units * p.num_n_blocks;
  const int num_kb = p.taps * p.cin_blocks;

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (elect_one()) {
      int stage = 0, stem_loaded = 0;
      uint32_t phase = 0;
      TileWalk walk(unit, num_units, num_tiles, p.tile_period);
      for (int tile; walk.next(tile);) {
        const int m_unit = (int)p.div_nb.div((uint32_t)tile), n_blk = tile - m_unit * p.num_n_blocks;
        const int m_blk = m_unit * (kPair ? 2 : 1) + (int)rank;
        int base_w = 0, base_h = 0, n_img = 0;
        if (p.a_im2col) {
          const int m0 = m_blk * BLOCK_M;
          n_img = (int)p.div_pq.div((uint32_t)m0);
          const int rem = m0 - n_img * p.P * p.Q;
          const int pp = (int)p.div_q.div((uint32_t)rem), qq = rem - pp * p.Q;
          base_w = qq * p.stride - p.pad;
          base_h = pp * p.stride - p.pad;
        }
        // k-blocks in (tap, channel-block) order, walked with counters instead of div/mod
        int r = 0, sx = 0, cb = 0;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty + stage, phase ^ 1);
          uint8_t* sA = smem + stage * L::STAGE_BYTES;
          uint8_t* sB = sA + A_STAGE_BYTES;
          if (kStem) {
            // stage kb = filter row kb of the 4x4 window: four taps, four loads.  kStages is a multiple of 4, so
            // ring slot s only ever holds filter row s % 4: its weight block is loaded on the slot's first use and
            // stays there for every later tile (the whole filter is 32 KB) — half the TMA operations per tile.
            static_assert(!kStem || L::kStages % 4 == 0, "stem: ring slots must map to fixed filter rows");
            const bool first_use = stem_loaded < kStages;
            mbar_arrive_expect_tx(full + stage, first_use ? L::STAGE_BYTES : A_STAGE_BYTES);
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              tma_load_im2col_4d(sA + t * (BLOCK_M * 32), &tmA, full + stage, 0, base_w, base_h, n_img, (uint16_t)t, (uint16_t)kb);
              if (first_use) tma_load_2d(sB + t * (BLOCK_N * 32), &tmB, full + stage, (kb * 4 + t) * 16, n_blk * BLOCK_N);
            }
            stem_loaded += first_use ? 1 : 0;
          } else if (kPair) {
            // both CTAs' bytes are credited to the leader's barrier; the peer only announces itself
            if (rank == 0) mbar_arrive_expect_tx(full + stage, 2 * L::STAGE_BYTES);
            else mbar_arrive_cluster(full + stage, 0);
            if (p.a_im2col)
              tma_load_im2col_4d_2sm(sA, &tmA, full + stage, cb * BLOCK_K, base_w, base_h, n_img, (uint16_t)sx, (uint16_t)r);
            else
              tma_load_2d_2sm(sA, &tmA, full + stage, kb * BLOCK_K, m_blk * BLOCK_M);
            tma_load_2d_2sm(sB, &tmB, full + stage, kb * BLOCK_K, n_blk * BLOCK_N + (int)rank * (BLOCK_N / 2));
          } else {
            mbar_arrive_expect_tx(full + stage, L::STAGE_BYTES);
            if (p.a_im2col) {
              tma_load_im2col_4d(sA, &tmA, full + stage, cb * BLOCK_K, base_w, base_h, n_img, (uint16_t)sx, (uint16_t)r);
            } else {
              tma_load_2d(sA, &tmA, full + stage, kb * BLOCK_K, m_blk * BLOCK_M);
            }
            tma_load_2d(sB, &tmB, full + stage, kb * BLOCK_K, n_blk * BLOCK_N);
          }
          if (++cb == p.cin_blocks) {
            cb = 0;
            if (++sx == p.S) sx = 0, ++r;
          }
          if (++stage == kStages) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ================================ MMA issuer (pair mode: leader CTA only) ======
    constexpr int kM = kPair ? 2 * BLOCK_M : BLOCK_M;
    constexpr uint32_t idesc = kTf32 ? umma_idesc_tf32(kM, BLOCK_N, 0, 0) : umma_idesc_f16(kFmt == 0 ? 1 : 0, kM, BLOCK_N, 0, 0);
    // K-major SW128: SBO = 8 rows x 128 B; one MMA advances 32 bytes along K.  Stem: K-major SW32, 8 rows x 32 B.
    constexpr uint64_t desc_base = kStem ? umma_desc_base(0, 256, kLayoutSw32) : umma_desc_base(0, 1024);
    constexpr uint32_t kStepA = kStem ? BLOCK_M * 32 : 32, kStepB = kStem ? BLOCK_N * 32 : 32;
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    TileWalk walk(unit, num_units, num_tiles, p.tile_period);
    for (int tile; walk.next(tile);) {
      mbar_wait(tempty + acc, acc_phase ^ 1);                       // epilogue has drained this accumulator
      fence_after_thread_sync();
      const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(full + stage, phase);
        fence_after_thread_sync();
        if (elect_one()) {
          const uint32_t a_addr = smem_u32(smem + stage * L::STAGE_BYTES);
          const uint32_t b_addr = a_addr + A_STAGE_BYTES;
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            const uint64_t da = umma_desc(desc_base, a_addr + k * kStepA), db = umma_desc(desc_base, b_addr + k * kStepB);
            if (kTf32) {
              if (kPair) umma_tf32_2sm(d_tmem, da, db, idesc, (kb | k) != 0);
              else umma_tf32(d_tmem, da, db, idesc, (kb | k) != 0);
            } else {
              if (kPair) umma_f16_2sm(d_tmem, da, db, idesc, (kb | k) != 0);
              else umma_f16(d_tmem, da, db, idesc, (kb | k) != 0);
            }
          }
          if (kPair) {
            umma_commit_2sm(empty + stage, 3);                      // frees the slot in BOTH CTAs
            if (kb == num_kb - 1) umma_commit_2sm(tfull + acc, 3);  // wakes both epilogues
          } else {
            umma_commit(empty + stage);                             // smem slot free once these MMAs retire
            if (kb == num_kb - 1) umma_commit(tfull + acc);         // accumulator complete
          }
        }
        __syncwarp();
        if (++stage == kStages) stage = 0, phase ^= 1;
      }
      if ((acc ^= 1) == 0) acc_phase ^= 1;
    }
  } else if (warp >= 2) {
    // ================================ epilogue ====================================
    const int quarter = warp & 3;                     // TMEM lanes [32*quarter, +32) are visible to this warp
    const int half = (warp - 2) >> 2;                 // the two warps of a quarter take alternate 32-column groups
    const int row = quarter * 32 + lane;              // row of the 128-row tile owned by this thread
    const int tid_e = threadIdx.x - 64;               // 0..255
    const bool leader = tid_e == 0;
    // The tile leaves in passes of at most 128 columns: a 256-wide tile re-uses a 32 KB staging buffer twice
    // instead of owning 64 KB, and the 32 KB saved buy one more pipeline stage — these kernels are bound by
    // TMA latency x bytes in flight (ncu: tensor pipe 30 % active, L2 and DRAM far from their limits).
    constexpr int PASS_N = L::PASS_N;
    constexpr int PASSES = BLOCK_N / PASS_N;          // 16-bit: 1, 1, 2; fp32: 1, 2, 4
    constexpr int BOX_N = L::BOX_N;                   // columns per 128-byte staging box row (64 / 32)
    constexpr int PAIRS = PASS_N / 2;                 // column pairs per pass for the statistics
    constexpr int G = kEpiThreads / PAIRS;            // row groups (8 or 4)
    constexpr int GPP = PASS_N / 64;                  // 32-column groups per warp and pass (1 or 2)
    const int pair = tid_e % PAIRS, rg = tid_e / PAIRS;
    const uint32_t out_stage_s = smem_u32(out_stage);
    // statistics: this thread owns column pair `pair` of every pass and rows rg, rg+G, ...; G divides 8, so the
    // swizzle term (row & 7) repeats every 8/G rows: its byte offsets are computed once, outside the row loop
    constexpr int SWZ = 8 / G;
    uint32_t st_off[SWZ];
    {
      const int col = pair * 2, byte = (col % BOX_N) * F::ES, chunk = byte >> 4;
#pragma unroll
      for (int i = 0; i < SWZ; ++i) {
        const int r = rg + i * G;
        st_off[i] = out_stage_s + (col / BOX_N) * 16384 + (byte & 15) + r * 128 + ((chunk ^ (r & 7)) << 4);
      }
    }
    float2 st_sum[PASSES], st_sq[PASSES];            // packed: add.f32x2 / fma.rn.f32x2 halve the FP32 issue slots
#pragma unroll
    for (int ps = 0; ps < PASSES; ++ps) st_sum[ps] = st_sq[ps] = make_float2(0.f, 0.f);
    int st_nblk = -1;
    float* const st_tgt = p.stats_rep ? p.stats_rep + (size_t)(blockIdx.x % kStatReplicas) * 2 * p.N : p.stats;
    auto flush_stats = [&]() {
#pragma unroll
      for (int ps = 0; ps < PASSES; ++ps) {
        const int c = st_nblk * BLOCK_N + ps * PASS_N + pair * 2;
        if (c < p.N) atomicAdd(st_tgt + c, st_sum[ps].x), atomicAdd(st_tgt + p.N + c, st_sq[ps].x);
        if (c + 1 < p.N) atomicAdd(st_tgt + c + 1, st_sum[ps].y), atomicAdd(st_tgt + p.N + c + 1, st_sq[ps].y);
        st_sum[ps] = st_sq[ps] = make_float2(0.f, 0.f);
      }
    };
    int acc = 0;
    uint32_t acc_phase = 0, obuf = 0;          // obuf: staging buffer of the current pass (kTwoOut: alternates)
    TileWalk walk(unit, num_units, num_tiles, p.tile_period);
    for (int tile; walk.next(tile);) {
      const int m_unit = (int)p.div_nb.div((uint32_t)tile), n_blk = tile - m_unit * p.num_n_blocks;
      const int m_blk = m_unit * (kPair ? 2 : 1) + (int)rank;
      const int n0 = n_blk * BLOCK_N;
      if (p.stats && st_nblk != n_blk) {
        if (st_nblk >= 0) flush_stats();
        st_nblk = n_blk;
      }
      mbar_wait(tfull + acc, acc_phase);
      fence_after_thread_sync();
      uint32_t v[GPP][32];
      auto issue_ld = [&](int ps) {
#pragma unroll
        for (int b = 0; b < GPP; ++b) {
          const int g = ps * (PASS_N / 32) + half + 2 * b;          // 32-column group of the tile
          tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * BLOCK_N + g * 32, v[b]);
        }
      };
      issue_ld(0);
#pragma unroll
      for (int ps = 0; ps < PASSES; ++ps) {
        tmem_ld_wait();
        if (ps == PASSES - 1) {
          // accumulator fully read: hand the TMEM stage back to the MMA warp
          fence_before_thread_sync();
          __syncwarp();
          if (lane == 0) {
            if (kPair) mbar_arrive_cluster(tempty + acc, 0);       // the MMA warp lives in the leader CTA
            else mbar_arrive(tempty + acc);
          }
        }
        // the store that last used this staging buffer must have finished reading it
        const uint32_t ob = kTwoOut ? obuf * L::OUT_BYTES : 0u;
        if (kTwoOut) obuf ^= 1u;
        if (leader) {
          if (kTwoOut) tma_store_wait_read1();
          else tma_store_wait_read0();
        }
        named_bar_sync(1, kEpiThreads);
#pragma unroll
        for (int b = 0; b < GPP; ++b) {
          const int lg = half + 2 * b;                               // group within the pass
          if (p.bias) {
#pragma unroll
            for (int c = 0; c < 32; ++c) {
              const int n = n0 + ps * PASS_N + lg * 32 + c;
              v[b][c] = __float_as_uint(__uint_as_float(v[b][c]) + (n < p.N ? __ldg(p.bias + n) : 0.f));
            }
          }
          if (kTf32) {
            // fp32 output: this warp's 32 columns are one full 128-byte row of box lg
            const uint32_t dst = out_stage_s + ob + lg * 16384 + row * 128;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              const uint4 w = make_uint4(v[b][4 * c], v[b][4 * c + 1], v[b][4 * c + 2], v[b][4 * c + 3]);
              sts128(dst + ((c ^ (row & 7)) << 4), w);
            }
          } else {
            const uint32_t dst = out_stage_s + ob + (lg >> 1) * 16384 + row * 128;
#pragma unroll
            for (int c = 0; c < 4; ++c) {               // four 16-byte chunks of this group's half row
              uint4 w;
              w.x = pack2<kFmt>(__uint_as_float(v[b][8 * c + 0]), __uint_as_float(v[b][8 * c + 1]));
              w.y = pack2<kFmt>(__uint_as_float(v[b][8 * c + 2]), __uint_as_float(v[b][8 * c + 3]));
              w.z = pack2<kFmt>(__uint_as_float(v[b][8 * c + 4]), __uint_as_float(v[b][8 * c + 5]));
              w.w = pack2<kFmt>(__uint_as_float(v[b][8 * c + 6]), __uint_as_float(v[b][8 * c + 7]));
              sts128(dst + ((((lg & 1) * 4 + c) ^ (row & 7)) << 4), w);
            }
          }
        }
        fence_proxy_async();                     // generic-proxy smem writes -> visible to the TMA engine
        named_bar_sync(1, kEpiThreads);
        if (leader) {
#pragma unroll
          for (int j = 0; j < PASS_N / BOX_N; ++j) {
            const int nc = n0 + ps * PASS_N + j * BOX_N;
            if (nc < p.N) {
              if (p.accumulate)
                tma_reduce_add_2d(&tmD, out_stage + ob + j * 16384, nc, m_blk * BLOCK_M);
              else
                tma_store_2d(&tmD, out_stage + ob + j * 16384, nc, m_blk * BLOCK_M);
            }
          }
          tma_store_commit();
        }
        // the next pass's accumulator columns travel TMEM -> registers while the statistics below are computed
        if (ps + 1 < PASSES) issue_ld(ps + 1);
        if (p.stats) {
          // per-channel sum / sum of squares of the ROUNDED outputs, read back from the staged columns
          // (rows past M are exact zeros: their A rows were zero-filled by TMA)
#pragma unroll 4
          for (int rb = 0; rb < BLOCK_M / 8; ++rb) {     // 8 rows per step: rows rg + 8*rb + i*G
#pragma unroll
            for (int i = 0; i < SWZ; ++i) {
              const float2 f = kTf32 ? lds64f(st_off[i] + ob + rb * (8 * 128)) : unpack2<kFmt>(lds32(st_off[i] + ob + rb * (8 * 128)));
              st_sum[ps] = __fadd2_rn(st_sum[ps], f);
              st_sq[ps] = __ffma2_rn(f, f, st_sq[ps]);
            }
          }
        }
      }
      if ((acc ^= 1) == 0) acc_phase ^= 1;
    }
    if (p.stats && st_nblk >= 0) flush_stats();
    if (p.stats && p.stats_rep) {
      // last CTA to get here folds the replicas (ticket lives right after them)
      __threadfence();
      named_bar_sync(1, kEpiThreads);
      volatile uint32_t* last_flag = tmem_slot + 1;
      if (leader) {
        unsigned int* ticket = reinterpret_cast<unsigned int*>(p.stats_rep + (size_t)kStatReplicas * 2 * p.N);
        *last_flag = atomicAdd(ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
      }
      named_bar_sync(1, kEpiThreads);
      if (*last_flag) {
        __threadfence();
        // one CTA alone adds 16 replicas x 2N floats: 128-bit loads, 16 of them in flight per thread (with scalar loads
        // this tail cost 10-17 us on the 1024- and 2048-channel layers, half as much as their whole main loop)
        if ((p.N & 1) == 0 && ((reinterpret_cast<uintptr_t>(p.stats) | reinterpret_cast<uintptr_t>(p.stats_rep)) & 15) == 0) {
          for (int c4 = tid_e; c4 < p.N / 2; c4 += kEpiThreads) {
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int r = 0; r < kStatReplicas; ++r) {
              const float4 v = __ldcg(reinterpret_cast<const float4*>(p.stats_rep + (size_t)r * 2 * p.N) + c4);
              a.x += v.x, a.y += v.y, a.z += v.z, a.w += v.w;
            }
            reinterpret_cast<float4*>(p.stats)[c4] = a;
          }
        } else {
          for (int c = tid_e; c < 2 * p.N; c += kEpiThreads) {
            float a = 0.f;
#pragma unroll
            for (int r = 0; r < kStatReplicas; ++r) a += __ldcg(p.stats_rep + (size_t)r * 2 * p.N + c);
            p.stats[c] = a;
          }
        }
        if (p.sync.world > 1) {
          // K6: the statistics all-reduce of SyncBatchNorm (torch:nn/modules/_functions.py:7-124 runs stats ->
          // all_gather -> combine as separate launches + a library collective) happens in the producing GEMM
          named_bar_sync(1, kEpiThreads);
          exchange_sum(p.sync, p.stats, p.stats, 2 * p.N, tid_e, kEpiThreads, [] { named_bar_sync(1, kEpiThreads); });
        }
      }
    }
    if (leader) tma_store_wait0();
  }
  fence_before_thread_sync();
  if (kPair) {
    cluster_sync_all();            // neither CTA may leave while the other can still signal its barriers / read its smem
    if (warp == 1) tmem_dealloc_2sm<2 * BLOCK_N>(tmem_base);
  } else {
    __syncthreads();
    if (warp == 1) tmem_dealloc<2 * BLOCK_N>(tmem_base);
  }
}

// ================================================================================================
// wgrad: dW[Cout, Ktot] += dY^T * im2col(X); both operands MN-major (pixels are the reduction axis).
struct WgradParams {
  int Mpix;                 // reduction length (N*P*Q)
  int Cout, Ktot;           // output is [Cout, Ktot] fp32, row pitch ldw
  int Kvalid;               // columns >= Kvalid are computed but never written
  int ldw;
  int num_m_blocks;         // ceil(Cout / 128)
  int num_n_blocks;         // ceil(Ktot / BLOCK_N)
  int num_kb;               // ceil(Mpix / pixels per k-block)
  int kb_per_split, num_splits;
  int cin_blocks, S;        // column chunk j (one MN atom wide) -> tap = j / cin_blocks, c0 = (j % cin_blocks) * atom
  int P, Q, stride, pad;
  int b_im2col;
  FastDiv div_pq, div_q, div_cin;
  float* dw;
};

// MN-major operands: a TMA box ("chunk") is KB_PIX pixel rows of 128 bytes, i.e. one MN atom of 64 16-bit or 32 fp32
// columns.  A k-block is 64 pixels (16-bit) or 32 pixels (fp32): the same 48 KB per stage for a 128 x 256 tile and
// four MMAs per stage in both formats.
template <int BLOCK_N, int ES = 2>
struct WgradSmem {
  static constexpr int MN_ATOM = 128 / ES;                                 // columns per chunk: 64 / 32
  static constexpr int KB_PIX = ES == 4 ? 32 : 64;                         // pixels per k-block
  static constexpr uint32_t CHUNK_BYTES = KB_PIX * 128;                    // 8 KB / 4 KB
  static constexpr int A_CHUNKS = BLOCK_M / MN_ATOM, B_CHUNKS = BLOCK_N / MN_ATOM;
  static constexpr uint32_t A_BYTES = A_CHUNKS * CHUNK_BYTES;              // 128 couts x KB_PIX pixels
  static constexpr uint32_t B_BYTES = B_CHUNKS * CHUNK_BYTES;              // BLOCK_N cols x KB_PIX pixels
  static constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int kStages = BLOCK_N == 256 ? 4 : 6;
  static constexpr uint32_t BAR_OFFSET = kStages * STAGE_BYTES;
  static constexpr uint32_t TOTAL = BAR_OFFSET + 256 + 1024;
};

// kStem: X is the space-to-depth stem input (16 channels per pixel): the 256 output columns are 16 taps x 16
// channels, loaded as sixteen [64 pixels x 32 B] blocks per stage in 32-byte-swizzled MN-major layout.
template <int BLOCK_N, int kFmt, bool kStem = false>
__global__ void __launch_bounds__(kNumThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmA /* dY [Mpix, Cout] */,
                  const __grid_constant__ CUtensorMap tmB /* X: tiled [Mpix, C] or im2col (C,W,H,N) */,
                  const WgradParams p) {
  using F = Fmt<kFmt>;
  constexpr bool kTf32 = kFmt == kFmtTf32;
  constexpr int UMMA_K = F::UMMA_K;
  static_assert(!(kStem && kTf32), "the space-to-depth stem kernels are 16-bit only");
  using L = WgradSmem<BLOCK_N, F::ES>;
  constexpr int MN_ATOM = L::MN_ATOM, KB_PIX = L::KB_PIX, A_CHUNKS = L::A_CHUNKS, B_CHUNKS = L::B_CHUNKS;
  constexpr uint32_t CHUNK_BYTES = L::CHUNK_BYTES;
  constexpr int kStages = L::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + L::BAR_OFFSET);
  uint64_t* empty = full + kStages;
  uint64_t* tfull = empty + kStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(full + i, 1);
      mbar_init(empty + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull + i, 1);
      mbar_init(tempty + i, kEpiThreads / 32);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<2 * BLOCK_N>(tmem_slot);
  fence_before_thread_sync();
  __syncthreads();
  fence_after_thread_sync();
  const uint32_t tmem_base = *tmem_slot;

  const int num_tiles = p.num_m_blocks * p.num_n_blocks;
  const int num_items = num_tiles * p.num_splits;

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        const int tile = item % num_tiles, split = item / num_tiles;
        const int m_blk = tile % p.num_m_blocks, n_blk = tile / p.num_m_blocks;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.num_kb);
        // the (tap, channel-block) of each 64-column chunk of this output tile is fixed for the whole item
        int ch_c[B_CHUNKS], ch_s[B_CHUNKS], ch_r[B_CHUNKS];
#pragma unroll
        for (int j = 0; j < B_CHUNKS; ++j) {
          const int chunk = n_blk * B_CHUNKS + j;                  // global column chunk index (MN_ATOM columns each)
          if (p.b_im2col) {
            const int tap = (int)p.div_cin.div((uint32_t)chunk);
            const int cb = chunk - tap * p.cin_blocks;
            const int r = tap / p.S;
            // chunks past Ktot: channel coordinate past C -> zero fill (tap clamped to stay in range)
            ch_c[j] = chunk * MN_ATOM < p.Ktot ? cb * MN_ATOM : (1 << 20);
            ch_s[j] = tap - r * p.S;
            ch_r[j] = min(r, 15);
          } else {
            ch_c[j] = chunk * MN_ATOM, ch_s[j] = 0, ch_r[j] = 0;
          }
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty + stage, phase ^ 1);
          uint8_t* sA = smem + stage * L::STAGE_BYTES;
          uint8_t* sB = sA + L::A_BYTES;
          mbar_arrive_expect_tx(full + stage, L::STAGE_BYTES);
          const int pix0 = kb * KB_PIX;
          // A: boxes of [KB_PIX pixels][MN_ATOM couts]; couts past Cout are zero-filled
#pragma unroll
          for (int j = 0; j < A_CHUNKS; ++j)
            tma_load_2d(sA + j * CHUNK_BYTES, &tmA, full + stage, m_blk * BLOCK_M + j * MN_ATOM, pix0);
          if (kStem) {
            const int n_img = (int)p.div_pq.div((uint32_t)pix0);
            const int rem = pix0 - n_img * p.P * p.Q;
            const int pp = (int)p.div_q.div((uint32_t)rem), qq = rem - pp * p.Q;
#pragma unroll
            for (int t = 0; t < 16; ++t)
              tma_load_im2col_4d(sB + t * 2048, &tmB, full + stage, 0, qq - p.pad, pp - p.pad, n_img, (uint16_t)(t & 3),
                                 (uint16_t)(t >> 2));
          } else if (p.b_im2col) {
            const int n_img = (int)p.div_pq.div((uint32_t)pix0);
            const int rem = pix0 - n_img * p.P * p.Q;
            const int pp = (int)p.div_q.div((uint32_t)rem), qq = rem - pp * p.Q;
            const int base_w = qq * p.stride - p.pad, base_h = pp * p.stride - p.pad;
#pragma unroll
            for (int j = 0; j < B_CHUNKS; ++j)
              tma_load_im2col_4d(sB + j * CHUNK_BYTES, &tmB, full + stage, ch_c[j], base_w, base_h, n_img,
                                 (uint16_t)ch_s[j], (uint16_t)ch_r[j]);
          } else {
#pragma unroll
            for (int j = 0; j < B_CHUNKS; ++j)
              tma_load_2d(sB + j * CHUNK_BYTES, &tmB, full + stage, ch_c[j], pix0);
          }
          if (++stage == kStages) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = kTf32 ? umma_idesc_tf32(BLOCK_M, BLOCK_N, 1, 1) : umma_idesc_f16(kFmt == 0 ? 1 : 0, BLOCK_M, BLOCK_N, 1, 1);
    // MN-major SW128: LBO = distance between MN atoms (one TMA box), SBO = 8 K-rows x 128 B.  32-bit operands only
    // exist in the layout swizzled in 32-byte units with a 4-row period: SBO = 4 K-rows x 128 B.
    constexpr uint64_t desc_base = kTf32 ? umma_desc_base(CHUNK_BYTES, 512, kLayoutSw128Base32) : umma_desc_base(CHUNK_BYTES, 1024);
    // stem B operand, MN-major SW32: LBO = one [64 pixels x 32 B] tap block, SBO = 8 pixel rows x 32 B
    constexpr uint64_t desc_base_b = kStem ? umma_desc_base(2048, 256, kLayoutSw32) : desc_base;
    constexpr uint32_t kStepB = kStem ? UMMA_K * 32 : UMMA_K * 128;
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int split = item / num_tiles;
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(kb0 + p.kb_per_split, p.num_kb);
      mbar_wait(tempty + acc, acc_phase ^ 1);
      fence_after_thread_sync();
      const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full + stage, phase);
        fence_after_thread_sync();
        if (elect_one()) {
          const uint32_t a_addr = smem_u32(smem + stage * L::STAGE_BYTES);
          const uint32_t b_addr = a_addr + L::A_BYTES;
#pragma unroll
          for (int k = 0; k < KB_PIX / UMMA_K; ++k) {   // UMMA_K (16 / 8) pixel rows x 128 B per MMA
            const uint64_t da = umma_desc(desc_base, a_addr + k * UMMA_K * 128), db = umma_desc(desc_base_b, b_addr + k * kStepB);
            if (kTf32) umma_tf32(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            else umma_f16(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(empty + stage);
          if (kb == kb1 - 1) umma_commit(tfull + acc);
        }
        __syncwarp();
        if (++stage == kStages) stage = 0, phase ^= 1;
      }
      if (kb1 > kb0) {
        if ((acc ^= 1) == 0) acc_phase ^= 1;
      }
    }
  } else {
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;                 // two warps per lane quarter, alternate 32-column groups
    const int row = quarter * 32 + lane;
    int acc = 0;
    uint32_t acc_phase = 0;
    const bool vec_ok = (p.ldw % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.dw) & 15) == 0);
    for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
      const int tile = item % num_tiles, split = item / num_tiles;
      const int kb0 = split * p.kb_per_split;
      if (kb0 >= p.num_kb) continue;
      const int m_blk = tile % p.num_m_blocks, n_blk = tile / p.num_m_blocks;
      mbar_wait(tfull + acc, acc_phase);
      fence_after_thread_sync();
      const int cout = m_blk * BLOCK_M + row;
#pragma unroll
      for (int j = half; j < BLOCK_N / 32; j += 2) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * BLOCK_N + j * 32, v);
        tmem_ld_wait();
        const int col0 = n_blk * BLOCK_N + j * 32;
        if (cout < p.Cout && col0 < p.Kvalid) {
          float* dst = p.dw + (size_t)cout * p.ldw + col0;
          if (vec_ok && col0 + 32 <= p.Kvalid) {
#pragma unroll
            for (int c = 0; c < 32; c += 4)
              red_add_v4(dst + c, __uint_as_float(v[c]), __uint_as_float(v[c + 1]), __uint_as_float(v[c + 2]),
                         __uint_as_float(v[c + 3]));
          } else {
#pragma unroll
            for (int c = 0; c < 32; ++c)
              if (col0 + c < p.Kvalid) atomicAdd(dst + c, __uint_as_float(v[c]));
          }
        }
      }
      fence_before_thread_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty + acc);
      if ((acc ^= 1) == 0) acc_phase ^= 1;
    }
  }
  fence_before_thread_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc<2 * BLOCK_N>(tmem_base);
}

// ================================================================================================
// Host side: tensor maps through the driver entry points (no link-time libcuda dependency).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);
static EncodeTiledFn g_tiled = nullptr;
static EncodeIm2colFn g_im2col = nullptr;

static int load_driver_fns() {
  if (g_tiled && g_im2col) return 0;
  cudaDriverEntryPointQueryResult q;
  void* f = nullptr;
  PDB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q));
  PDB_CHECK_ARG(f && q == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled not available in this driver");
  g_tiled = (EncodeTiledFn)f;
  PDB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &f, cudaEnableDefault, &q));
  PDB_CHECK_ARG(f && q == cudaDriverEntryPointSuccess, "cuTensorMapEncodeIm2col not available in this driver");
  g_im2col = (EncodeIm2colFn)f;
  return 0;
}

// `operand`: the map feeds an MMA operand.  fp32 operands are loaded as TFLOAT32: the copy engine rounds them to the
// 10-bit mantissa (round to nearest) on their way into shared memory, where the tensor core would otherwise just
// ignore the low 13 bits (truncation, a systematic bias) — the same conversion cuDNN's TF32 kernels apply.
static CUtensorMapDataType tm_dtype(int dtype, bool operand) {
  if (dtype == PDB_FP32) return operand ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  return dtype == PDB_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
}
static inline int elem_bytes(int dtype) { return dtype == PDB_FP32 ? 4 : 2; }

// 2-D row-major [rows, cols] matrix with row pitch ld (elements); box = [box_rows, box_cols]
static int make_tiled_2d(CUtensorMap* m, const void* ptr, int dtype, uint64_t rows, uint64_t cols, uint64_t ld,
                         uint32_t box_rows, uint32_t box_cols, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B,
                         bool operand = true) {
  const uint64_t es = (uint64_t)elem_bytes(dtype);
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * es};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  PDB_CHECK_ARG((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (ld * es) % 16 == 0,
                "tensor map: base must be 16-byte aligned and the row pitch a multiple of 16 bytes (ld=%llu)",
                (unsigned long long)ld);
  CUresult r = g_tiled(m, tm_dtype(dtype, operand), 2, const_cast<void*>(ptr), dims, strides, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PDB_CHECK_ARG(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with %d (rows=%llu cols=%llu ld=%llu box=%ux%u)",
                (int)r, (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_rows, box_cols);
  return 0;
}

// NHWC activation as (C, W, H, N) in im2col mode; `pixels` consecutive output positions per load.
// pad_hi < 0: symmetric padding.  cbox / swz: channels per load and the matching shared-memory swizzle (64 / 128B for
// the regular layers, 16 / 32B for the space-to-depth stem).
static int make_im2col_4d(CUtensorMap* m, const void* ptr, int dtype, int N, int H, int W, int C, int R, int S,
                          int stride, int pad, uint32_t pixels, int pad_hi = -1, uint32_t cbox = 0,
                          CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  const cuuint64_t es = (cuuint64_t)elem_bytes(dtype);
  if (cbox == 0) cbox = (uint32_t)(128 / es);        // one 128-byte swizzle row of channels
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * es, (cuuint64_t)W * C * es, (cuuint64_t)H * W * C * es};
  if (pad_hi < 0) pad_hi = pad;
  int lower[2] = {-pad, -pad};
  int upper[2] = {pad_hi - (S - 1), pad_hi - (R - 1)};
  cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  PDB_CHECK_ARG((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (C * es) % 16 == 0,
                "im2col tensor map: base must be 16-byte aligned and the pixel pitch a multiple of 16 bytes");
  CUresult r = g_im2col(m, tm_dtype(dtype, true), 4, const_cast<void*>(ptr), dims, strides, lower, upper, cbox, pixels, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PDB_CHECK_ARG(r == CUDA_SUCCESS, "cuTensorMapEncodeIm2col failed with %d (N=%d H=%d W=%d C=%d R=%d S=%d stride=%d pad=%d)",
                (int)r, N, H, W, C, R, S, stride, pad);
  return 0;
}

template <typename K>
static int set_smem(K kernel, uint32_t bytes) {
  PDB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return 0;
}

static int walk_period(int units, int n_blocks) {
  int a = units, b = n_blocks;
  while (b) {
    const int t = a % b;
    a = b, b = t;
  }
  return n_blocks / a;          // n_blocks / gcd(units, n_blocks)
}

template <int kFmt, bool kTwoOut = false>
static int launch_fprop_pair(const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& d, FpropParams p,
                             cudaStream_t st) {
  constexpr int BN = 256;
  static bool configured[64] = {};
  auto kernel = conv_fprop_kernel<BN, kFmt, true, false, kTwoOut>;
  int dev = 0;
  PDB_CUDA(cudaGetDevice(&dev));
  const uint32_t smem = FpropSmem<BN, true, Fmt<kFmt>::ES, kTwoOut>::TOTAL;
  if (!configured[dev & 63]) {
    if (int e = set_smem(kernel, smem)) return e;
    configured[dev & 63] = true;
  }
  const int pair_tiles = ((p.num_m_blocks + 1) / 2) * p.num_n_blocks;
  int pairs = num_sms() / 2;
  if (pair_tiles < pairs) pairs = pair_tiles;
  p.tile_period = walk_period(pairs, p.num_n_blocks);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(kNumThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr, cfg.numAttrs = 1;
  PDB_CUDA(cudaLaunchKernelEx(&cfg, kernel, a, b, d, p));
  return 0;
}

template <int BLOCK_N, int kFmt, bool kStem = false, bool kTwoOut = false>
static int launch_fprop(const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& d, FpropParams p,
                        cudaStream_t st) {
  // the opt-in to > 48 KB of dynamic shared memory is a per-device function attribute (the single-process
  // DataParallel mode launches from one process on every GPU)
  static bool configured[64] = {};
  auto kernel = conv_fprop_kernel<BLOCK_N, kFmt, false, kStem, kTwoOut>;
  constexpr uint32_t smem = FpropSmem<BLOCK_N, false, Fmt<kFmt>::ES, kTwoOut>::TOTAL;
  int dev = 0;
  PDB_CUDA(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    if (int e = set_smem(kernel, smem)) return e;
    configured[dev & 63] = true;
  }
  const int tiles = p.num_m_blocks * p.num_n_blocks;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  p.tile_period = walk_period(grid, p.num_n_blocks);
  kernel<<<grid, kNumThreads, smem, st>>>(a, b, d, p);
  PDB_LAUNCH_CHECK();
  return 0;
}

template <int BLOCK_N, int kFmt, bool kStem = false>
static int launch_wgrad(const CUtensorMap& a, const CUtensorMap& b, const WgradParams& p, cudaStream_t st) {
  static bool configured[64] = {};
  auto kernel = conv_wgrad_kernel<BLOCK_N, kFmt, kStem>;
  constexpr uint32_t smem = WgradSmem<BLOCK_N, Fmt<kFmt>::ES>::TOTAL;
  int dev = 0;
  PDB_CUDA(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    if (int e = set_smem(kernel, smem)) return e;
    configured[dev & 63] = true;
  }
  const int items = p.num_m_blocks * p.num_n_blocks * p.num_splits;
  const int grid = items < num_sms() ? items : num_sms();
  kernel<<<grid, kNumThreads, smem, st>>>(a, b, p);
  PDB_LAUNCH_CHECK();
  return 0;
}

}  // namespace pdb
using namespace pdb;

extern "C" size_t pdb_conv_stats_scratch_floats(int K) { return (size_t)kStatReplicas * 2 * K + 32; }

namespace pdb {
int fill_sync(SyncPeers* sp, void* const* staging, void* const* pads, int rank, int world, int slot_stride, int row);   // collectives.cu
static int set_sync(FpropParams* p, const pdb_sync_peers* sync, const char* who) {
  p->sync = SyncPeers{};
  if (!sync || sync->world <= 1) return 0;
  PDB_CHECK_ARG(p->stats && p->stats_rep, "%s: the cross-GPU statistics exchange needs stats and stats_scratch", who);
  PDB_CHECK_ARG(2 * p->N <= sync->slot_stride, "%s: 2K=%d exceeds the staging slot (%d floats)", who, 2 * p->N, sync->slot_stride);
  return fill_sync(&p->sync, sync->staging, sync->pads, sync->rank, sync->world, sync->slot_stride, sync->row);
}
}  // namespace pdb

extern "C" int pdb_conv_fprop(const void* x, const void* w, void* y, int dtype, const pdb_conv_shape* s, int64_t lda,
                              int64_t ldd, float* stats, float* stats_scratch, const float* bias, int block_n,
                              int accumulate, const pdb_sync_peers* sync, void* stream) {
  PDB_CHECK_ARG(x && w && y && s, "pdb_conv_fprop: null pointer");
  PDB_CHECK_ARG(dtype == PDB_BF16 || dtype == PDB_FP16 || dtype == PDB_FP32,
                "pdb_conv_fprop: dtype must be PDB_BF16, PDB_FP16 or PDB_FP32");
  if (int e = load_driver_fns()) return e;
  const int es = elem_bytes(dtype), block_k = 128 / es;     // elements per k-block
  const int P = (s->H + 2 * s->pad - s->R) / s->stride + 1, Q = (s->W + 2 * s->pad - s->S) / s->stride + 1;
  const int64_t M = (int64_t)s->N * P * Q;
  const bool plain = s->R == 1 && s->S == 1 && s->stride == 1 && s->pad == 0;
  PDB_CHECK_ARG(M > 0 && M < (1ll << 31) && s->K > 0 && s->C > 0, "pdb_conv_fprop: bad shape");
  PDB_CHECK_ARG((s->C * es) % 16 == 0, "pdb_conv_fprop: rows of C=%d elements are not a multiple of 16 bytes", s->C);
  PDB_CHECK_ARG(plain || s->C % block_k == 0, "pdb_conv_fprop: im2col path needs C %% %d == 0 (C=%d)", block_k, s->C);
  if (lda <= 0) lda = s->C;
  if (ldd <= 0) ldd = s->K;
  int bn = block_n;
  bool pair = bn == 512;                  // 512: a 256-wide tile computed by a CTA pair (cta_group::2, M = 256)
  if (pair) bn = 256;
  if (bn <= 0) {
    bn = s->K <= 64 ? 64 : (s->K <= 128 ? 128 : 256);
    // Measured on the ResNet-50 shapes (profiles/conv_tune_pair.json): the pair halves each CTA's weight traffic and
    // wins 3-20 % once every CTA has six or more tiles to pipeline; on two-to-three-tile grids (7x7 maps) the coupled
    // pipelines of the two CTAs lose 6-11 % to the single-CTA tile.
    const int64_t tiles = ((M + BLOCK_M - 1) / BLOCK_M) * ((s->K + 255) / 256);
    pair = bn == 256 && tiles >= 6ll * num_sms();
  }
  PDB_CHECK_ARG(bn == 64 || bn == 128 || bn == 256, "pdb_conv_fprop: block_n must be 64, 128, 256 or 512 (pair)");

  FpropParams p{};
  p.M = (int)M, p.N = s->K;
  p.num_m_blocks = (int)((M + BLOCK_M - 1) / BLOCK_M);
  p.num_n_blocks = (s->K + bn - 1) / bn;
  p.taps = s->R * s->S, p.cin_blocks = (s->C + block_k - 1) / block_k, p.S = s->S;
  p.P = P, p.Q = Q, p.stride = s->stride, p.pad = s->pad;
  p.a_im2col = plain ? 0 : 1;
  p.accumulate = accumulate ? 1 : 0;
  p.div_pq.init((uint32_t)(P * Q)), p.div_q.init((uint32_t)Q), p.div_nb.init((uint32_t)p.num_n_blocks);
  p.stats = stats, p.stats_rep = stats ? stats_scratch : nullptr, p.bias = bias;
  if (int e = set_sync(&p, sync, "pdb_conv_fprop")) return e;

  CUtensorMap tmA, tmB, tmD;
  if (plain) {
    if (int e = make_tiled_2d(&tmA, x, dtype, (uint64_t)M, (uint64_t)s->C, (uint64_t)lda, BLOCK_M, block_k)) return e;
  } else {
    if (int e = make_im2col_4d(&tmA, x, dtype, s->N, s->H, s->W, s->C, s->R, s->S, s->stride, s->pad, BLOCK_M)) return e;
  }
  const uint64_t ktot = (uint64_t)p.taps * s->C;
  if (int e = make_tiled_2d(&tmB, w, dtype, (uint64_t)s->K, ktot, ktot, (uint32_t)(pair ? bn / 2 : bn), block_k)) return e;
  if (int e = make_tiled_2d(&tmD, y, dtype, (uint64_t)M, (uint64_t)s->K, (uint64_t)ldd, BLOCK_M, block_k,
                            CU_TENSOR_MAP_SWIZZLE_128B, false)) return e;

  cudaStream_t st = (cudaStream_t)stream;
  // Two alternating staging buffers (store of pass i overlaps the conversion of pass i+1, paid for with one pipeline
  // stage): measured on all 22 shapes in both precisions, -4..-6 % on two short-K layers (128->512, 256->1024) and
  // +2..+17 % on everything with a longer reduction — the store wait is not what bounds the epilogue.  Off unless
  // PDB200_FPROP_TWO_OUT=1 (kept for A/B timing and because it is tested).
  static const int two_out_env = [] { const char* e = getenv("PDB200_FPROP_TWO_OUT"); return e ? atoi(e) : 0; }();
  const bool two_out = two_out_env != 0 && bn >= 128;
  if (pair) {
    if (two_out)
      return dtype == PDB_BF16   ? launch_fprop_pair<0, true>(tmA, tmB, tmD, p, st)
             : dtype == PDB_FP16 ? launch_fprop_pair<1, true>(tmA, tmB, tmD, p, st)
                                 : launch_fprop_pair<2, true>(tmA, tmB, tmD, p, st);
    return dtype == PDB_BF16   ? launch_fprop_pair<0>(tmA, tmB, tmD, p, st)
           : dtype == PDB_FP16 ? launch_fprop_pair<1>(tmA, tmB, tmD, p, st)
                               : launch_fprop_pair<2>(tmA, tmB, tmD, p, st);
  }
#define GO(BN, TWO)                                                                  \
  (dtype == PDB_BF16   ? launch_fprop<BN, 0, false, TWO>(tmA, tmB, tmD, p, st)       \
   : dtype == PDB_FP16 ? launch_fprop<BN, 1, false, TWO>(tmA, tmB, tmD, p, st)       \
                       : launch_fprop<BN, 2, false, TWO>(tmA, tmB, tmD, p, st))
  if (bn == 64) return GO(64, false);
  if (bn == 128) return two_out ? GO(128, true) : GO(128, false);
  return two_out ? GO(256, true) : GO(256, false);
#undef GO
}

extern "C" int pdb_conv_wgrad(const void* x, const void* dy, float* dw, int dtype, const pdb_conv_shape* s, int64_t ldx,
                              int64_t lddy, int64_t ldw, int64_t kvalid, int splits, int block_n, void* stream) {
  PDB_CHECK_ARG(x && dy && dw && s, "pdb_conv_wgrad: null pointer");
  PDB_CHECK_ARG(dtype == PDB_BF16 || dtype == PDB_FP16 || dtype == PDB_FP32,
                "pdb_conv_wgrad: dtype must be PDB_BF16, PDB_FP16 or PDB_FP32");
  if (int e = load_driver_fns()) return e;
  const int es = elem_bytes(dtype);
  const int atom = 128 / es;                    // MN atom: columns per TMA box (64 / 32)
  const int kb_pix = es == 4 ? 32 : 64;         // pixels per k-block
  const CUtensorMapSwizzle mn_swz = es == 4 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  const int P = (s->H + 2 * s->pad - s->R) / s->stride + 1, Q = (s->W + 2 * s->pad - s->S) / s->stride + 1;
  const int64_t M = (int64_t)s->N * P * Q;
  const bool plain = s->R == 1 && s->S == 1 && s->stride == 1 && s->pad == 0;
  PDB_CHECK_ARG(M > 0 && M < (1ll << 31) && s->K > 0 && s->C > 0, "pdb_conv_wgrad: bad shape");
  PDB_CHECK_ARG((s->C * es) % 16 == 0 && (s->K * es) % 16 == 0, "pdb_conv_wgrad: rows of C / K elements must be multiples of 16 bytes");
  PDB_CHECK_ARG(plain || s->C % atom == 0, "pdb_conv_wgrad: im2col path needs C %% %d == 0 (C=%d)", atom, s->C);
  if (ldx <= 0) ldx = s->C;
  if (lddy <= 0) lddy = s->K;
  const int ktot = s->R * s->S * s->C;
  if (ldw <= 0) ldw = ktot;
  int bn = block_n;
  if (bn <= 0) bn = ktot <= 64 ? 64 : (ktot <= 128 ? 128 : 256);
  PDB_CHECK_ARG(bn == 64 || bn == 128 || bn == 256, "pdb_conv_wgrad: block_n must be 64, 128 or 256");

  WgradParams p{};
  p.Mpix = (int)M, p.Cout = s->K, p.Ktot = ktot, p.ldw = (int)ldw;
  p.Kvalid = kvalid > 0 && kvalid < ktot ? (int)kvalid : ktot;
  p.num_m_blocks = (s->K + BLOCK_M - 1) / BLOCK_M;
  p.num_n_blocks = (ktot + bn - 1) / bn;
  p.num_kb = (int)((M + kb_pix - 1) / kb_pix);
  const int tiles = p.num_m_blocks * p.num_n_blocks;
  if (splits <= 0) {
    // One wave of persistent CTAs: items = tiles * splits must not spill over the SM count (a second,
    // nearly empty wave doubles the time — measured with bench_tools/conv_tune.py), and every extra split
    // costs a full tile of fp32 reductions; keep at least 4 k-blocks of work per item.
    splits = num_sms() / tiles;
    const int max_splits = (p.num_kb * kb_pix / 64 + 3) / 4;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
  }
  p.kb_per_split = (p.num_kb + splits - 1) / splits;
  p.num_splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
  p.cin_blocks = (s->C + atom - 1) / atom, p.S = s->S;
  p.P = P, p.Q = Q, p.stride = s->stride, p.pad = s->pad;
  p.b_im2col = plain ? 0 : 1;
  p.div_pq.init((uint32_t)(P * Q)), p.div_q.init((uint32_t)Q), p.div_cin.init((uint32_t)p.cin_blocks);
  p.dw = dw;

  CUtensorMap tmA, tmB;
  if (int e = make_tiled_2d(&tmA, dy, dtype, (uint64_t)M, (uint64_t)s->K, (uint64_t)lddy, kb_pix, atom, mn_swz)) return e;
  if (plain) {
    if (int e = make_tiled_2d(&tmB, x, dtype, (uint64_t)M, (uint64_t)s->C, (uint64_t)ldx, kb_pix, atom, mn_swz)) return e;
  } else {
    if (int e = make_im2col_4d(&tmB, x, dtype, s->N, s->H, s->W, s->C, s->R, s->S, s->stride, s->pad, kb_pix, -1, atom, mn_swz))
      return e;
  }
  cudaStream_t st = (cudaStream_t)stream;
#define GO(BN)                                                     \
  (dtype == PDB_BF16   ? launch_wgrad<BN, 0>(tmA, tmB, p, st)      \
   : dtype == PDB_FP16 ? launch_wgrad<BN, 1>(tmA, tmB, p, st)      \
                       : launch_wgrad<BN, 2>(tmA, tmB, p, st))
  if (bn == 64) return GO(64);
  if (bn == 128) return GO(128);
  return GO(256);
#undef GO
}

// ---- the 7x7 / stride-2 / pad-3 stem on the space-to-depth input ---------------------------------------------------
// xs: [N, Hs, Ws, 16] 16-bit (Hs = H/2, Ws = W/2; channel (dy*2 + dx)*3 + c holds x[c][2h+dy][2w+dx], 12..15 zero), so
// the stem becomes a 4x4 / stride-1 convolution with padding 2 before and 1 after: output (p, q) reads rows
// p-2 .. p+1.  ws: [64, 256] 16-bit, column (r*4 + s)*16 + ch (pdb_stem_weight_s2d).  No im2col matrix exists: TMA
// im2col mode gathers the taps straight from xs (ref delegate: cuDNN's 7x7 stem on NCHW input).
extern "C" int pdb_stem_fprop(const void* xs, const void* ws, void* y, int dtype, int N, int Hs, int Ws, float* stats,
                              float* stats_scratch, const pdb_sync_peers* sync, void* stream) {
  PDB_CHECK_ARG(xs && ws && y && N > 0 && Hs > 0 && Ws > 0, "pdb_stem_fprop: bad arguments");
  PDB_CHECK_ARG(dtype == PDB_BF16 || dtype == PDB_FP16, "pdb_stem_fprop: dtype must be PDB_BF16 or PDB_FP16");
  if (int e = load_driver_fns()) return e;
  const int64_t M = (int64_t)N * Hs * Ws;
  PDB_CHECK_ARG(M < (1ll << 31), "pdb_stem_fprop: too many pixels");
  FpropParams p{};
  p.M = (int)M, p.N = 64;
  p.num_m_blocks = (int)((M + BLOCK_M - 1) / BLOCK_M), p.num_n_blocks = 1;
  p.taps = 4, p.cin_blocks = 1, p.S = 4;         // four k-blocks: one filter row each
  p.P = Hs, p.Q = Ws, p.stride = 1, p.pad = 2;
  p.a_im2col = 1, p.accumulate = 0;
  p.div_pq.init((uint32_t)(Hs * Ws)), p.div_q.init((uint32_t)Ws), p.div_nb.init(1u);
  p.stats = stats, p.stats_rep = stats ? stats_scratch : nullptr, p.bias = nullptr;
  if (int e = set_sync(&p, sync, "pdb_stem_fprop")) return e;
  CUtensorMap tmA, tmB, tmD;
  if (int e = make_im2col_4d(&tmA, xs, dtype, N, Hs, Ws, 16, 4, 4, 1, 2, BLOCK_M, 1, 16, CU_TENSOR_MAP_SWIZZLE_32B)) return e;
  if (int e = make_tiled_2d(&tmB, ws, dtype, 64, 256, 256, 64, 16, CU_TENSOR_MAP_SWIZZLE_32B)) return e;
  if (int e = make_tiled_2d(&tmD, y, dtype, (uint64_t)M, 64, 64, BLOCK_M, 64)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == PDB_BF16 ? launch_fprop<64, 0, true>(tmA, tmB, tmD, p, st)
                           : launch_fprop<64, 1, true>(tmA, tmB, tmD, p, st);
}

// dws[64, 256] (fp32, same column order as ws) += dY[M, 64]^T * im2col(xs); unpack with pdb_stem_dw_unpack.
extern "C" int pdb_stem_wgrad(const void* xs, const void* dy, float* dws, int dtype, int N, int Hs, int Ws, int splits,
                              void* stream) {
  PDB_CHECK_ARG(xs && dy && dws && N > 0 && Hs > 0 && Ws > 0, "pdb_stem_wgrad: bad arguments");
  PDB_CHECK_ARG(dtype == PDB_BF16 || dtype == PDB_FP16, "pdb_stem_wgrad: dtype must be PDB_BF16 or PDB_FP16");
  if (int e = load_driver_fns()) return e;
  const int64_t M = (int64_t)N * Hs * Ws;
  PDB_CHECK_ARG(M < (1ll << 31), "pdb_stem_wgrad: too many pixels");
  WgradParams p{};
  p.Mpix = (int)M, p.Cout = 64, p.Ktot = 256, p.Kvalid = 256, p.ldw = 256;
  p.num_m_blocks = 1, p.num_n_blocks = 1;
  p.num_kb = (int)((M + 63) / 64);
  if (splits <= 0) splits = num_sms();
  if (splits > p.num_kb) splits = p.num_kb;
  p.kb_per_split = (p.num_kb + splits - 1) / splits;
  p.num_splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
  p.cin_blocks = 1, p.S = 4;
  p.P = Hs, p.Q = Ws, p.stride = 1, p.pad = 2;
  p.b_im2col = 1;
  p.div_pq.init((uint32_t)(Hs * Ws)), p.div_q.init((uint32_t)Ws), p.div_cin.init(1u);
  p.dw = dws;
  CUtensorMap tmA, tmB;
  if (int e = make_tiled_2d(&tmA, dy, dtype, (uint64_t)M, 64, 64, 64, 64)) return e;
  if (int e = make_im2col_4d(&tmB, xs, dtype, N, Hs, Ws, 16, 4, 4, 1, 2, 64, 1, 16, CU_TENSOR_MAP_SWIZZLE_32B)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  return dtype == PDB_BF16 ? launch_wgrad<256, 0, true>(tmA, tmB, p, st) : launch_wgrad<256, 1, true>(tmA, tmB, p, st);
}
