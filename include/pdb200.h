/* pdb200.h — C ABI of the sm_100a kernel library of pytorch_distributed_b200 (lib/libpdb200.so).
 *
 * Conventions
 *   - every entry point returns 0 on success, non-zero on failure; pdb_last_error() then gives a
 *     human-readable message (thread-local);
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing synchronises;
 *   - activations are NHWC ("channels last"); `dtype` selects their storage format: PDB_BF16 / PDB_FP16
 *     (16-bit, the amp mode) or PDB_FP32 (fp32 storage; the GEMMs then run tcgen05 kind::tf32, which is what
 *     the reference's three fp32 scripts get from cuDNN's TF32 default).  "16" in an entry point's name is
 *     historical: unless a comment says otherwise every activation kernel takes all three tags.  Statistics,
 *     master weights, gradients and the loss are always fp32;
 *   - "ref:" cites the reference script that causes the delegate kernel to run, "torch:" the
 *     delegate whose behaviour is reproduced (see SURVEY.md §2.4 for the K-numbers).
 */
#ifndef PDB200_H_
#define PDB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PDB_BF16 0
#define PDB_FP16 1
#define PDB_FP32 2

/* Peers of a small cross-GPU sum that a kernel performs itself (SyncBatchNorm statistics in the convolution epilogue and
 * in the BatchNorm backward reduction; pdb_small_allreduce uses the same protocol).  staging[r]: rank r's symmetric
 * staging area float[2][world][slot_stride] as mapped in THIS process; pads[r]: rank r's signal pad, of which row `row`
 * (world 32-bit slots) is used.  Exchanges alternate between the two halves of the staging area, so one meeting per
 * exchange suffices; all ranks must issue the same sequence of exchanges on a row.  world <= 1 or a null pointer: no
 * exchange.  (K6; torch:nn/modules/_functions.py:7-209 does this with all_gather / all_reduce calls per BN layer.) */
typedef struct {
  void* staging[16];
  void* pads[16];
  int rank, world;
  int slot_stride;
  int row;
} pdb_sync_peers;

const char* pdb_last_error(void);
int pdb_version(void);

/* ---------------------------------------------------------------------------------------------
 * K9 — fused SGD step over flat arenas (ref:restnet_ddp.py:30 optimizer.step();
 * torch:optim/sgd.py:322-380).  g' = (inv_scale? g*inv_scale : g) + wd*p;
 * buf = first_step ? g' : momentum*buf + (1-dampening)*g';  p -= lr * (nesterov ? g'+momentum*buf : buf).
 * If found_inf (device float) is non-null and non-zero the update is skipped on the device (no host
 * sync; torch:amp/grad_scaler.py:363-373 does a .item()).  Optionally writes the 16-bit weight
 * shadow and zeroes the gradient.  buf may be null when momentum == 0. */
int pdb_sgd_step(float* p, float* g, float* buf, void* shadow16, int shadow_dtype, int64_t n, float lr,
                 float momentum, float dampening, float weight_decay, int nesterov, int maximize,
                 int first_step, const float* inv_scale, const float* found_inf, int zero_grad,
                 void* stream);

/* K13 — non-finite check of a flat fp32 range: *found_inf = 1.0f if any element is inf/nan
 * (torch:amp/grad_scaler.py:349-361 _amp_foreach_non_finite_check_and_unscale_; the un-scale itself
 * is folded into pdb_sgd_step / the all-reduce epilogue). */
int pdb_check_finite(const float* g, int64_t n, float* found_inf, void* stream);
/* amp scale update on the device (one thread): scale *= backoff and tracker = 0 if *found_inf > 0; otherwise
 * ++tracker and, when it reaches `interval`, scale *= growth (kept if that would overflow) and tracker = 0.
 * Also writes *inv_scale = 1/scale and clears *found_inf.  Delegate: torch GradScaler.update / amp_update_scale. */
int pdb_scaler_update(float* scale, int* growth_tracker, float* found_inf, float* inv_scale, float growth,
                      float backoff, int interval, void* stream);

/* fp32 -> 16-bit copy of a flat range (weight shadow refresh after load_state_dict). */
int pdb_cast_f32_to_16(const float* src, void* dst, int dtype, int64_t n, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K10 — fused cross-entropy forward + backward (ref:restnet_ddp.py:28; torch:nn/functional.py:3421).
 * logits [rows, classes] in any activation dtype; labels int64.  loss_sum += sum_i -log softmax(x_i)[y_i] / rows
 * (caller zeroes *loss beforehand); dlogits (nullable) = (softmax - onehot) * grad_scale / rows, stored as
 * dlogits_dtype: the logits' dtype or PDB_FP32 (what the Python layer uses, so that the amp loss scale is applied
 * before the gradient is rounded to 16 bits).  grad_scale may be null (1.0) or a device float (amp loss scale). */
int pdb_xent_fwd_bwd(const void* logits, int dtype, const int64_t* labels, int rows, int classes,
                     float* loss, void* dlogits, int dlogits_dtype, const float* grad_scale, void* stream);

/* K14 — top-k hit counters (ref:restnet_ddp.py:58-60).  out[j] += #rows whose label ranks < ks[j].
 * Ties are broken by index like torch.topk(sorted=True). */
int pdb_topk_correct(const void* logits, int dtype, const int64_t* labels, int rows, int classes,
                     const int* ks, int nk, float* out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K8 — BatchNorm (+ReLU, +residual) on NHWC 16-bit activations, fp32 statistics
 * (torch:nn/modules/batchnorm.py:144-210; eps 1e-5, momentum 0.1, unbiased running variance). */

/* Per-channel sum and sum of squares of x[rows, C] into sums[2*C] (sum | sumsq).  Two-stage:
 * per-block partials go to `scratch` (>= pdb_bn_stats_scratch_bytes(C) bytes, any contents) and the
 * last block to finish folds them, in a fixed order, into sums — deterministic, no atomics on
 * the data path.  The convolution epilogue normally produces these sums itself; this entry point is
 * for layers whose producer is not one of our GEMMs and for tests. */
size_t pdb_bn_stats_scratch_bytes(int C);
int pdb_bn_stats(const void* x, int dtype, int64_t rows, int C, float* sums, void* scratch,
                 size_t scratch_bytes, void* stream);

/* K6 — SyncBatchNorm statistics (torch:nn/modules/_functions.py:7-124): pdb_bn_stats with the cross-GPU
 * reduction inside the kernel.  The block that folds the partials pushes the [2C] sums into slot `rank` of
 * every peer's symmetric staging buffer (float[2][world][slot_stride], P2P stores), meets the peers on the
 * staging allocation's signal pad and adds the `world` slots: `sums` comes back global.  One rank per GPU. */
int pdb_bn_stats_sync(const void* x, int dtype, int64_t rows, int C, float* sums, void* scratch,
                      size_t scratch_bytes, void* const* staging, void* const* pads, int rank, int world,
                      int slot_stride, void* stream);

/* Same result as pdb_bn_stats for callers that cannot obtain device memory for the scratch buffer
 * (hosts whose GPU memory is fully committed to their own pools, code in the middle of a step where
 * cudaMalloc/cudaFree would synchronise the device, embedders that may not call the CUDA allocator).
 * This entry point has no counterpart in the reference.  The partial sums go to a fixed ~512 KB
 * array inside the library's device image, resident on every GPU that loads the library until exit.
 * Limit: C <= 512 (a multiple of 8); larger C returns an error.  The first call on each device
 * looks up the array's device address (do it before any stream capture); calls that overlap on one
 * GPU share the array and race — run them one after another.  Allocates nothing. */
int pdb_bn_stats_noalloc(const void* x, int dtype, int64_t rows, int C, float* sums, void* stream);

/* y = [relu]( (x - mean) * rstd * gamma + beta [+ residual] ).  mean/rstd are derived from
 * sums[2*C] over `count` elements per channel (count may exceed rows under SyncBN) inside the
 * kernel; block 0 stores save_mean/save_rstd and updates running_mean/var/num_batches_tracked when
 * given.  relu_mask (nullable, rows*C/8 bytes): bit k of byte i says whether element 8*i+k passed the
 * ReLU — the backward kernels read it instead of the 16x larger activation. */
int pdb_bn_apply(const void* x, const void* residual, void* y, int dtype, int64_t rows, int C,
                 const float* sums, double count, const float* gamma, const float* beta, float eps,
                 float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked,
                 float* save_mean, float* save_rstd, int relu, uint8_t* relu_mask, void* stream);

/* Inference-mode BN: uses running statistics. */
int pdb_bn_apply_eval(const void* x, const void* residual, void* y, int dtype, int64_t rows, int C,
                      const float* running_mean, const float* running_var, const float* gamma,
                      const float* beta, float eps, int relu, void* stream);

/* Backward reductions: with g = dz masked by the ReLU (relu_mask bits if given, else z > 0 if z is
 * given, else unmasked), dsums[0:C] += sum g, dsums[C:2C] += sum g * (x - mean) * rstd.  dsums must be
 * zeroed by the caller.
 * scratch (nullable): pdb_bn_bwd_reduce_scratch_floats(C) ZEROED floats; when given, the per-block partials are
 * spread over 16 replicas and folded by the last block (dsums is then overwritten, not accumulated) — this
 * removes the atomic contention that otherwise costs ~45 us per launch.
 * sync (nullable; needs scratch): SyncBatchNorm — the block that folds the replicas also sums dsums over the peer GPUs. */
size_t pdb_bn_bwd_reduce_scratch_floats(int C);
int pdb_bn_bwd_reduce(const void* dz, const void* z, const uint8_t* relu_mask, const void* x, int dtype,
                      int64_t rows, int C, const float* mean, const float* rstd, float* dsums, float* scratch,
                      const pdb_sync_peers* sync, void* stream);

/* The same reduction without atomics: per-block partials in `scratch` (>= pdb_bn_stats_scratch_bytes(C) bytes, ticket
 * word zero) folded in block order by the last block — bit-reproducible from run to run.  Used by the engine's
 * deterministic mode (PDB200_DETERMINISTIC=1), whose "same input, same bits" property is the step's race detector. */
int pdb_bn_bwd_reduce_det(const void* dz, const void* z, const uint8_t* relu_mask, const void* x, int dtype,
                          int64_t rows, int C, const float* mean, const float* rstd, float* dsums, void* scratch,
                          size_t scratch_bytes, void* stream);

/* dx = gamma*rstd * (g - dsums[c]/count - xhat * dsums[C+c]/count);  optionally also writes
 * dres = g (gradient of the residual branch).  Block 0 accumulates dgamma += dsums[C:], dbeta +=
 * dsums[:C] (fp32 gradient arena views; may be null). */
int pdb_bn_bwd_apply(const void* dz, const void* z, const uint8_t* relu_mask, const void* x, void* dx, void* dres,
                     int dtype, int64_t rows, int C, const float* mean, const float* rstd, const float* gamma,
                     const float* dsums, double count, float* dgamma, float* dbeta, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K12 — pooling on NHWC 16-bit. */
int pdb_maxpool3x3s2_fwd(const void* x, void* y, uint8_t* argmax, int dtype, int N, int H, int W, int C,
                         void* stream);
int pdb_maxpool3x3s2_bwd(const void* dy, const uint8_t* argmax, void* dx, int dtype, int N, int H, int W,
                         int C, void* stream);
int pdb_avgpool_fwd(const void* x, void* y, int dtype, int N, int HW, int C, void* stream);
int pdb_avgpool_bwd(const void* dy, void* dx, int dtype, int N, int HW, int C, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Layout / elementwise helpers. */
/* NCHW fp32 -> NHWC 16-bit (the loader yields NCHW fp32, ref:restnet_ddp.py:25). */
int pdb_nchw_f32_to_nhwc16(const float* x, void* y, int dtype, int N, int C, int H, int W, void* stream);
/* Stem im2col: NCHW fp32 [N,3,H,W] -> [N*Ho*Wo, Kpad] 16-bit patches for the 7x7/2 pad-3 stem conv
 * (K = 147, columns ordered (kh, kw, c), zero padded to Kpad). */
int pdb_stem_im2col(const float* x, void* col, int dtype, int N, int H, int W, int Kpad, void* stream);
/* The 7x7 / stride-2 / pad-3 stem WITHOUT an im2col matrix: a 4x4 / stride-1 convolution over the space-to-depth
 * input.  pdb_stem_s2d: x [N,3,H,W] fp32 -> xs [N,H/2,W/2,16] 16-bit, channel (dy*2+dx)*3+c = x[c][2h+dy][2w+dx],
 * 12..15 zero.  pdb_stem_weight_s2d: w [K,7,7,3] 16-bit -> ws [K,256], column (r*4+s)*16 + channel.
 * pdb_stem_fprop: y [N*Hs*Ws, 64] = conv (+ BatchNorm sums like pdb_conv_fprop).  pdb_stem_wgrad: dws [64,256] fp32
 * += dY^T im2col(xs) (zero it first), then pdb_stem_dw_unpack folds it into dw [K,7,7,3] (accumulate != 0: add). */
int pdb_stem_s2d(const float* x, void* xs, int dtype, int N, int H, int W, void* stream);
int pdb_stem_weight_s2d(const void* w, void* ws, int dtype, int K, void* stream);
int pdb_stem_fprop(const void* xs, const void* ws, void* y, int dtype, int N, int Hs, int Ws, float* stats,
                   float* stats_scratch, const pdb_sync_peers* sync, void* stream);
int pdb_stem_wgrad(const void* xs, const void* dy, float* dws, int dtype, int N, int Hs, int Ws, int splits,
                   void* stream);
int pdb_stem_dw_unpack(const float* dws, float* dw, int K, int accumulate, void* stream);
/* out = a + b (16-bit), used where a residual sum is not fused into a producer. */
int pdb_add16(const void* a, const void* b, void* out, int dtype, int64_t n, void* stream);
/* dst[n, 2h, 2w, :] = src[n, h, w, :], zeros elsewhere: zero-upsample by 2 (stride-2 dgrad input). */
int pdb_upsample2_zero(const void* src, void* dst, int dtype, int N, int Ho, int Wo, int C, void* stream);
/* out[n,h,w,:] = a[n,h,w,:] + (h,w even ? b[n,h/2,w/2,:] : 0): merges the stride-2 1x1 dgrad. */
int pdb_add_strided2(const void* a, const void* b, void* out, int dtype, int N, int H, int W, int C,
                     void* stream);

/* ---------------------------------------------------------------------------------------------
 * K7 / K11 — convolutions and the fc layer as tcgen05 implicit GEMMs (ref:restnet_ddp.py:26,29
 * model(samples) / loss.backward(); delegates: cuDNN 9.22 conv fwd/dgrad/wgrad, cuBLASLt linear). */
typedef struct {
  int N, H, W, C;     /* input activation, NHWC; a plain GEMM is H = W = 1, N = rows */
  int K;              /* output channels */
  int R, S;           /* filter height / width */
  int stride, pad;
} pdb_conv_shape;

/* y[N*P*Q, K] = conv(x, w).  w is [K, R, S, C] 16-bit (K-major rows of length R*S*C).  lda / ldd:
 * row pitch in elements of x (plain 1x1 path only) and y; <= 0 means dense.  stats (nullable):
 * fp32 [2K], per-channel sum | sum of squares of the rounded outputs are ADDED to it (zero it
 * first) — the BatchNorm statistics pass fused into the epilogue.  bias (nullable): fp32 [K].
 * block_n: 64 / 128 / 256, or 0 for the default heuristic.  accumulate != 0: y += conv(x, w) through a TMA
 * reduction store (the residual-branch gradient is summed in the dgrad epilogue instead of by a separate
 * elementwise kernel).  dgrad is the same call on dY with the
 * weights rearranged by pdb_weight_dgrad_layout. */
/* stats_scratch (nullable): pdb_conv_stats_scratch_floats(K) ZEROED floats.  When given, the CTAs spread their
 * per-channel partial sums over 16 replicas and the last CTA folds them (stats is then overwritten): without it
 * every CTA adds to the same 2K addresses and the L2 atomic units serialise (20-30 us on wide layers). */
size_t pdb_conv_stats_scratch_floats(int K);
/* sync (nullable; needs stats + stats_scratch): SyncBatchNorm — the CTA that folds the replicas also sums the [2K]
 * statistics over the peer GPUs before the kernel ends (no separate collective launch). */
int pdb_conv_fprop(const void* x, const void* w, void* y, int dtype, const pdb_conv_shape* s, int64_t lda,
                   int64_t ldd, float* stats, float* stats_scratch, const float* bias, int block_n, int accumulate,
                   const pdb_sync_peers* sync, void* stream);

/* dw[K, R*S*C] (fp32, row pitch ldw) += dy[N*P*Q, K]^T * im2col(x).  The pixel range is split over
 * CTAs (`splits`, 0 = heuristic) and partial tiles are added with fp32 vector reductions, so dw
 * must hold the running gradient (zero for a fresh step).  Only columns < kvalid are written
 * (<= 0: all R*S*C). */
int pdb_conv_wgrad(const void* x, const void* dy, float* dw, int dtype, const pdb_conv_shape* s, int64_t ldx,
                   int64_t lddy, int64_t ldw, int64_t kvalid, int splits, int block_n, void* stream);

/* wt[C, R, S, K] = w[K, R-1-r, S-1-s, C]: weight layout for running dgrad through pdb_conv_fprop. */
int pdb_weight_dgrad_layout(const void* w, void* wt, int dtype, int K, int R, int S, int C, void* stream);
/* The same for n convolutions in one launch: table_dev is a device array of n x {w_off, wt_off, K, R, S,
 * C} (int64, element offsets into w_base / wt_base). */
int pdb_weight_dgrad_layout_batched(const void* w_base, void* wt_base, const int64_t* table_dev, int n, int dtype,
                                    void* stream);
/* dst[rows, cols_dst] = src[rows, cols_src] zero-padded on the right (16-bit). */
int pdb_pad_rows16(const void* src, void* dst, int dtype, int rows, int cols_src, int cols_dst, void* stream);
/* out[c] += sum_r a[r, c] (16-bit in, fp32 out): fc bias gradient. */
int pdb_colsum16(const void* a, float* out, int dtype, int rows, int cols, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K1-K6 — collectives over NVLink peer memory (csrc/comm/collectives.cu).  `bufs[r]` / `pads[r]` are
 * rank r's symmetric buffer base and signal pad as mapped in THIS process (torch symmetric memory
 * provides allocation and handle exchange only).  Every rank must call with the same arguments and the
 * same `blocks` (<= pdb_comm_max_blocks); kernels of different ranks wait for each other, so each rank
 * needs its own GPU.  Spins are bounded (20 s) and trap. */
int pdb_comm_max_blocks(int world, size_t signal_pad_bytes);
/* Profiling aid: buf = device array of 3 x capacity uint64.  While set, launch i of pdb_allreduce_inplace
 * stamps buf[3i..3i+2] with the GPU's global timer (ns) at kernel entry, after the first cross-rank meeting, and at exit:
 * inside a training step the first interval is the wait for the slowest rank, the second the collective itself.
 * buf = NULL switches it off.  Returns how many launches were traced into the previously set buffer. */
int pdb_comm_set_trace(void* buf, int capacity);

/* K1 (ref:restnet_ddp.py:29 loss.backward() -> Reducer all_reduce_bucket -> ncclAllReduce): in-place
 * all-reduce of fp32 elements [start, start+n) of every rank's buffer: reduce-scatter by P2P loads (or
 * one multimem.ld_reduce per 16 B when `multicast` is the NVLS address of the same buffer), multiply
 * by scale (1/world) and *inv_scale (device float, nullable: amp un-scale), raise flags[r][0] = 1 on
 * every rank if the result is non-finite (nullable), all-gather by P2P stores / multimem.st. */
int pdb_allreduce_inplace(void* const* bufs, void* const* pads, void* const* flags, void* multicast, int rank,
                          int world, int64_t start, int64_t n, float scale, const float* inv_scale, int blocks,
                          void* stream);
/* K1 with a 16-bit wire format (the fp16_compress / bf16_compress comm hooks, torch:distributed/
 * algorithms/ddp_comm_hooks/default_hooks.py:57-134): gradients stay fp32 in the arena, each rank packs
 * its range (x pre_scale = 1/world) into the first half of a symmetric 16-bit staging buffer, shard owners
 * sum the peers' packed data in fp32, apply *inv_scale and the non-finite check, store the 16-bit result
 * into the second half of every rank's staging buffer, and every rank unpacks it back to fp32.
 * staging[r] holds 2 * staging_half_elems 16-bit elements. */
int pdb_allreduce_inplace_wire16(void* const* bufs, void* const* pads, void* const* flags, void* const* staging,
                                 int64_t staging_half_elems, int wire_dtype, int rank, int world, int64_t start,
                                 int64_t n, float pre_scale, const float* inv_scale, int blocks, void* stream);
/* K2/K3 (DDP _sync_module_states / per-iteration buffer broadcast): every rank != root pulls
 * [start, start+n) from root's buffer. */
int pdb_broadcast_inplace(void* const* bufs, void* const* pads, int rank, int world, int root, int64_t start,
                          int64_t n, int blocks, void* stream);
/* K1 for small buckets (up to 64 * 8192 floats; the Python layer switches at 256 KB): one-shot variant of
 * pdb_allreduce_inplace — meet, every rank reads all copies and adds them in rank order (bit-identical sums on every
 * rank), meet, write the local copy.  Same fused scale / un-scale / non-finite check; flags[rank][0] is raised locally. */
int pdb_allreduce_oneshot(void* const* bufs, void* const* pads, void* const* flags, int rank, int world, int64_t start,
                          int64_t n, float scale, const float* inv_scale, void* stream);
/* K5/K6 (ref:restnet_ddp.py:63-64 metric reduce; SyncBatchNorm statistics): one-shot in-place sum over
 * ranks of data[0:n] through a symmetric staging area float[2][world][slot_stride] (see pdb_sync_peers). */
int pdb_small_allreduce(void* const* staging, void* const* pads, int rank, int world, float* data, int n,
                        int slot_stride, void* stream);
/* K4a-K4c (ref:resnet_dp.py:21 DataParallel scatter / replicate / gather): the current GPU pulls
 * `bytes` from another GPU's memory (single process, peer access is enabled on demand). */
int pdb_peer_pull(void* dst, const void* src, int64_t bytes, void* stream);
/* K4d (ref:resnet_dp.py:24 backward -> ReduceAddCoalesced -> nccl.reduce): root[0:n] += sum of the
 * peers' arrays, read over NVLink by the root GPU. */
int pdb_reduce_to_root(float* root, void* const* peers, int n_peers, int64_t n, void* stream);

/* K4b/K4d, second half of the SHARDED replicate / reduce of the DataParallel mode: for every q < n except `skip`, bytes
 * [q*shard_bytes, min((q+1)*shard_bytes, total_bytes)) of dst are copied from the same range of srcs[q] (the GPU of dst
 * reads its peers).  With it, parameter replicate = "replica r pulls shard r from GPU 0, then everybody gathers the
 * other shards from each other" and gradient reduce = "GPU r sums shard r of all replicas, then GPU 0 gathers the
 * shards": GPU 0's links carry ~2/n of an arena per direction instead of n-1 arenas (ref:resnet_dp.py:21-25,82;
 * delegates torch:nn/parallel/comm.py:50-168 broadcast_coalesced / reduce_add_coalesced). */
int pdb_peer_gather(void* dst, void* const* srcs, int n, int64_t shard_bytes, int64_t total_bytes, int skip,
                    void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PDB200_H_ */
