"""Tensor-level wrappers of the sm_100a kernels.  All launches go to the current CUDA stream."""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import check, count_launch, ptr, require, stream_ptr
from ._sigs import ConvShape, SyncPeersC

__all__ = ["dtype_tag", "sgd_step", "check_finite", "scaler_update", "cast_to_16", "xent_fwd_bwd", "topk_correct", "bn_stats",
           "bn_apply", "bn_apply_eval", "bn_bwd_reduce", "bn_bwd_reduce_det", "bn_bwd_reduce_scratch_floats", "bn_bwd_apply", "maxpool_fwd", "maxpool_bwd",
           "avgpool_fwd", "avgpool_bwd", "nchw_to_nhwc16", "stem_im2col", "stem_s2d", "stem_weight_s2d", "stem_fprop",
           "stem_wgrad", "stem_dw_unpack", "add16", "upsample2_zero",
           "add_strided2", "conv_fprop", "conv_wgrad", "weight_dgrad_layout", "weight_dgrad_layout_batched", "pad_rows16", "colsum16",
           "peer_copy", "peer_copy_", "peer_gather_", "reduce_to_root", "ConvShape", "SyncPeersC"]

_TAG = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}


def dtype_tag(dt: torch.dtype) -> int:
    try:
        return _TAG[dt]
    except KeyError:
        raise TypeError(f"native kernels take bf16/fp16/fp32 activations, got {dt}") from None


def _st(t: torch.Tensor):
    return stream_ptr(t.device)


def sgd_step(p, g, buf, shadow, *, lr, momentum, dampening, weight_decay, nesterov, maximize, first_step,
             inv_scale=None, found_inf=None, zero_grad=False) -> None:
    L = require()
    with torch.cuda.device(p.device):
        check(L.pdb_sgd_step(ptr(p), ptr(g), ptr(buf), ptr(shadow), dtype_tag(shadow.dtype) if shadow is not None else 0,
                             p.numel(), lr, momentum, dampening, weight_decay, int(nesterov), int(maximize),
                             int(first_step), ptr(inv_scale), ptr(found_inf), int(zero_grad), _st(p)), "pdb_sgd_step")
    count_launch()


def scaler_update(scale, tracker, found_inf, inv_scale, growth: float, backoff: float, interval: int) -> None:
    """GradScaler.update() as one device thread (K13): no tensor-select chain, no host sync."""
    assert scale.dtype == torch.float32 and tracker.dtype == torch.int32
    check(require().pdb_scaler_update(ptr(scale), ptr(tracker), ptr(found_inf), ptr(inv_scale), growth, backoff,
                                      interval, _st(scale)), "pdb_scaler_update")
    count_launch()


def check_finite(g: torch.Tensor, found_inf: torch.Tensor) -> None:
    check(require().pdb_check_finite(ptr(g), g.numel(), ptr(found_inf), _st(g)), "pdb_check_finite")
    count_launch()


def cast_to_16(src: torch.Tensor, dst: torch.Tensor) -> None:
    check(require().pdb_cast_f32_to_16(ptr(src), ptr(dst), dtype_tag(dst.dtype), src.numel(), _st(src)),
          "pdb_cast_f32_to_16")
    count_launch()


def xent_fwd_bwd(logits: torch.Tensor, labels: torch.Tensor, grad_scale: Optional[torch.Tensor] = None,
                 need_grad: bool = True, grad_dtype: Optional[torch.dtype] = torch.float32
                 ) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Mean cross-entropy and its gradient.  The gradient comes back in fp32 unless ``grad_dtype`` asks for the
    logits' own dtype: the amp loss scale must be applied before the rounding to 16 bits, not after."""
    logits = logits.contiguous()
    rows, classes = logits.shape
    loss = torch.zeros((), dtype=torch.float32, device=logits.device)
    dl = torch.empty(logits.shape, dtype=grad_dtype or logits.dtype, device=logits.device) if need_grad else None
    check(require().pdb_xent_fwd_bwd(ptr(logits), dtype_tag(logits.dtype), ptr(labels), rows, classes, ptr(loss),
                                     ptr(dl), dtype_tag(dl.dtype) if dl is not None else 0, ptr(grad_scale),
                                     _st(logits)), "pdb_xent_fwd_bwd")
    count_launch()
    return loss, dl


_ks_cache = {}


def topk_correct(logits: torch.Tensor, labels: torch.Tensor, ks: Sequence[int] = (1, 5)) -> torch.Tensor:
    logits = logits.contiguous()
    if logits.dtype not in _TAG:
        logits = logits.float()
    key = (tuple(ks), logits.device)
    if key not in _ks_cache:
        _ks_cache[key] = torch.tensor(list(ks), dtype=torch.int32, device=logits.device)
    out = torch.zeros(len(ks), dtype=torch.float32, device=logits.device)
    check(require().pdb_topk_correct(ptr(logits), dtype_tag(logits.dtype), ptr(labels), logits.shape[0],
                                     logits.shape[1], ptr(_ks_cache[key]), len(ks), ptr(out), _st(logits)),
          "pdb_topk_correct")
    count_launch()
    return out


_scratch = {}


def bn_stats_scratch(C: int, device) -> torch.Tensor:
    n = require().pdb_bn_stats_scratch_bytes(C)
    key = (torch.device(device), torch.cuda.current_stream(device).cuda_stream)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < n:
        buf = torch.zeros(n, dtype=torch.uint8, device=device)      # ticket must start at 0
        _scratch[key] = buf
    return buf


def bn_stats(x2d: torch.Tensor, sums: torch.Tensor, scratch: Optional[torch.Tensor] = None) -> None:
    rows, C = x2d.shape
    scratch = scratch if scratch is not None else bn_stats_scratch(C, x2d.device)
    check(require().pdb_bn_stats(ptr(x2d), dtype_tag(x2d.dtype), rows, C, ptr(sums), ptr(scratch), scratch.numel(),
                                 _st(x2d)), "pdb_bn_stats")
    count_launch()


def bn_apply(x, residual, y, sums, count, gamma, beta, eps, momentum, running_mean, running_var, nbt, save_mean,
             save_rstd, relu: bool, relu_mask=None) -> None:
    rows, C = x.shape
    check(require().pdb_bn_apply(ptr(x), ptr(residual), ptr(y), dtype_tag(x.dtype), rows, C, ptr(sums), float(count),
                                 ptr(gamma), ptr(beta), eps, momentum, ptr(running_mean), ptr(running_var), ptr(nbt),
                                 ptr(save_mean), ptr(save_rstd), int(relu), ptr(relu_mask), _st(x)), "pdb_bn_apply")
    count_launch()


def bn_apply_eval(x, residual, y, running_mean, running_var, gamma, beta, eps, relu: bool) -> None:
    rows, C = x.shape
    check(require().pdb_bn_apply_eval(ptr(x), ptr(residual), ptr(y), dtype_tag(x.dtype), rows, C, ptr(running_mean),
                                      ptr(running_var), ptr(gamma), ptr(beta), eps, int(relu), _st(x)),
          "pdb_bn_apply_eval")
    count_launch()


def bn_bwd_reduce_scratch_floats(C: int) -> int:
    return int(require().pdb_bn_bwd_reduce_scratch_floats(C))


def _sync_ref(sync):
    """ctypes reference to a SyncPeersC (or NULL): the in-kernel cross-GPU sum of SyncBatchNorm."""
    return ctypes.byref(sync) if sync is not None else None


def bn_bwd_reduce(dz, z, x, mean, rstd, dsums, relu_mask=None, scratch=None, sync=None) -> None:
    rows, C = x.shape
    check(require().pdb_bn_bwd_reduce(ptr(dz), ptr(z), ptr(relu_mask), ptr(x), dtype_tag(x.dtype), rows, C, ptr(mean),
                                      ptr(rstd), ptr(dsums), ptr(scratch), _sync_ref(sync), _st(x)), "pdb_bn_bwd_reduce")
    count_launch()


def bn_bwd_reduce_det(dz, z, x, mean, rstd, dsums, relu_mask=None, scratch=None) -> None:
    """Atomic-free, bit-reproducible variant (deterministic mode)."""
    rows, C = x.shape
    scratch = scratch if scratch is not None else bn_stats_scratch(C, x.device)
    check(require().pdb_bn_bwd_reduce_det(ptr(dz), ptr(z), ptr(relu_mask), ptr(x), dtype_tag(x.dtype), rows, C, ptr(mean),
                                          ptr(rstd), ptr(dsums), ptr(scratch), scratch.numel(), _st(x)),
          "pdb_bn_bwd_reduce_det")
    count_launch()


def bn_bwd_apply(dz, z, x, dx, dres, mean, rstd, gamma, dsums, count, dgamma, dbeta, relu_mask=None) -> None:
    rows, C = x.shape
    check(require().pdb_bn_bwd_apply(ptr(dz), ptr(z), ptr(relu_mask), ptr(x), ptr(dx), ptr(dres), dtype_tag(x.dtype), rows, C,
                                     ptr(mean), ptr(rstd), ptr(gamma), ptr(dsums), float(count), ptr(dgamma),
                                     ptr(dbeta), _st(x)), "pdb_bn_bwd_apply")
    count_launch()


def maxpool_fwd(x, y, argmax) -> None:
    N, H, W, C = x.shape
    check(require().pdb_maxpool3x3s2_fwd(ptr(x), ptr(y), ptr(argmax), dtype_tag(x.dtype), N, H, W, C, _st(x)),
          "pdb_maxpool3x3s2_fwd")
    count_launch()


def maxpool_bwd(dy, argmax, dx) -> None:
    N, H, W, C = dx.shape
    check(require().pdb_maxpool3x3s2_bwd(ptr(dy), ptr(argmax), ptr(dx), dtype_tag(dx.dtype), N, H, W, C, _st(dx)),
          "pdb_maxpool3x3s2_bwd")
    count_launch()


def avgpool_fwd(x, y) -> None:
    N, H, W, C = x.shape
    check(require().pdb_avgpool_fwd(ptr(x), ptr(y), dtype_tag(x.dtype), N, H * W, C, _st(x)), "pdb_avgpool_fwd")
    count_launch()


def avgpool_bwd(dy, dx) -> None:
    N, H, W, C = dx.shape
    check(require().pdb_avgpool_bwd(ptr(dy), ptr(dx), dtype_tag(dx.dtype), N, H * W, C, _st(dx)), "pdb_avgpool_bwd")
    count_launch()


def nchw_to_nhwc16(x: torch.Tensor, y: torch.Tensor) -> None:
    N, C, H, W = x.shape
    check(require().pdb_nchw_f32_to_nhwc16(ptr(x), ptr(y), dtype_tag(y.dtype), N, C, H, W, _st(x)),
          "pdb_nchw_f32_to_nhwc16")
    count_launch()


def stem_im2col(x: torch.Tensor, col: torch.Tensor) -> None:
    N, C, H, W = x.shape
    assert C == 3 and x.dtype == torch.float32 and x.is_contiguous()
    check(require().pdb_stem_im2col(ptr(x), ptr(col), dtype_tag(col.dtype), N, H, W, col.shape[-1], _st(x)),
          "pdb_stem_im2col")
    count_launch()


def stem_s2d(x: torch.Tensor, xs: torch.Tensor) -> None:
    """x [N,3,H,W] fp32 -> xs [N,H/2,W/2,16] 16-bit space-to-depth (K7 stem without an im2col matrix)."""
    N, C, H, W = x.shape
    assert C == 3 and x.dtype == torch.float32 and x.is_contiguous() and xs.shape == (N, H // 2, W // 2, 16)
    check(require().pdb_stem_s2d(ptr(x), ptr(xs), dtype_tag(xs.dtype), N, H, W, _st(x)), "pdb_stem_s2d")
    count_launch()


def stem_weight_s2d(w16: torch.Tensor, ws: torch.Tensor) -> None:
    """w16: [K, 7, 7, 3] (physical order) 16-bit -> ws [K, 256]."""
    K = ws.shape[0]
    assert w16.numel() == K * 147 and ws.shape == (K, 256) and w16.dtype == ws.dtype
    check(require().pdb_stem_weight_s2d(ptr(w16), ptr(ws), dtype_tag(ws.dtype), K, _st(ws)), "pdb_stem_weight_s2d")
    count_launch()


def stem_fprop(xs: torch.Tensor, ws: torch.Tensor, y: torch.Tensor, stats=None, stats_scratch=None, sync=None) -> None:
    N, Hs, Ws, C = xs.shape
    assert C == 16 and ws.shape == (64, 256) and y.numel() == N * Hs * Ws * 64
    check(require().pdb_stem_fprop(ptr(xs), ptr(ws), ptr(y), dtype_tag(xs.dtype), N, Hs, Ws, ptr(stats),
                                   ptr(stats_scratch), _sync_ref(sync), _st(xs)), "pdb_stem_fprop")
    count_launch()


def stem_wgrad(xs: torch.Tensor, dy: torch.Tensor, dws: torch.Tensor, splits: int = 0) -> None:
    """dws [64, 256] fp32 += dY^T im2col(xs); the caller zeroes dws."""
    N, Hs, Ws, C = xs.shape
    assert C == 16 and dws.shape == (64, 256) and dws.dtype == torch.float32 and dy.numel() == N * Hs * Ws * 64
    check(require().pdb_stem_wgrad(ptr(xs), ptr(dy), ptr(dws), dtype_tag(xs.dtype), N, Hs, Ws, splits, _st(xs)),
          "pdb_stem_wgrad")
    count_launch()


def stem_dw_unpack(dws: torch.Tensor, dw: torch.Tensor, accumulate: bool = True) -> None:
    K = dws.shape[0]
    assert dw.numel() == K * 147 and dw.dtype == torch.float32
    check(require().pdb_stem_dw_unpack(ptr(dws), ptr(dw), K, int(accumulate), _st(dws)), "pdb_stem_dw_unpack")
    count_launch()


def add16(a, b, out) -> None:
    check(require().pdb_add16(ptr(a), ptr(b), ptr(out), dtype_tag(a.dtype), a.numel(), _st(a)), "pdb_add16")
    count_launch()


def upsample2_zero(src, dst) -> None:
    N, Ho, Wo, C = src.shape
    check(require().pdb_upsample2_zero(ptr(src), ptr(dst), dtype_tag(src.dtype), N, Ho, Wo, C, _st(src)),
          "pdb_upsample2_zero")
    count_launch()


def add_strided2(a, b, out) -> None:
    N, H, W, C = a.shape
    check(require().pdb_add_strided2(ptr(a), ptr(b), ptr(out), dtype_tag(a.dtype), N, H, W, C, _st(a)),
          "pdb_add_strided2")
    count_launch()


def conv_fprop(x, w, y, shape: ConvShape, *, lda: int = 0, ldd: int = 0, stats=None, stats_scratch=None, bias=None,
               block_n: int = 0, accumulate: bool = False, sync=None) -> None:
    check(require().pdb_conv_fprop(ptr(x), ptr(w), ptr(y), dtype_tag(x.dtype), ctypes.byref(shape), lda, ldd,
                                   ptr(stats), ptr(stats_scratch), ptr(bias), block_n, int(accumulate), _sync_ref(sync),
                                   _st(x)), "pdb_conv_fprop")
    count_launch()


def conv_wgrad(x, dy, dw, shape: ConvShape, *, ldx: int = 0, lddy: int = 0, ldw: int = 0, kvalid: int = 0,
               splits: int = 0, block_n: int = 0) -> None:
    check(require().pdb_conv_wgrad(ptr(x), ptr(dy), ptr(dw), dtype_tag(x.dtype), ctypes.byref(shape), ldx, lddy, ldw,
                                   kvalid, splits, block_n, _st(x)), "pdb_conv_wgrad")
    count_launch()


def weight_dgrad_layout(w, wt, K, R, S, C) -> None:
    check(require().pdb_weight_dgrad_layout(ptr(w), ptr(wt), dtype_tag(w.dtype), K, R, S, C, _st(w)),
          "pdb_weight_dgrad_layout")
    count_launch()


def weight_dgrad_layout_batched(w_base, wt_base, table_dev, n: int) -> None:
    check(require().pdb_weight_dgrad_layout_batched(ptr(w_base), ptr(wt_base), ptr(table_dev), n,
                                                    dtype_tag(w_base.dtype), _st(w_base)),
          "pdb_weight_dgrad_layout_batched")
    count_launch()


def pad_rows16(src, dst, rows, cols_src, cols_dst) -> None:
    check(require().pdb_pad_rows16(ptr(src), ptr(dst), dtype_tag(src.dtype), rows, cols_src, cols_dst, _st(src)),
          "pdb_pad_rows16")
    count_launch()


def colsum16(a, out) -> None:
    rows, cols = a.shape
    check(require().pdb_colsum16(ptr(a), ptr(out), dtype_tag(a.dtype), rows, cols, _st(a)), "pdb_colsum16")
    count_launch()


# ---- peer-memory helpers (single-process DP mode); kernels land in csrc/comm/p2p.cu ----------------
def _has(name: str) -> bool:
    return hasattr(require(), name)


def peer_copy_(dst: torch.Tensor, src: torch.Tensor) -> None:
    """dst (on its own GPU) pulls src's bytes over NVLink (K4a/K4b/K4c)."""
    nbytes = dst.numel() * dst.element_size()
    if not _has("pdb_peer_pull") or nbytes % 16 or (dst.data_ptr() | src.data_ptr()) % 16:
        dst.copy_(src, non_blocking=True)       # odd sizes (e.g. a handful of labels): plain copy engine
        return
    assert dst.is_contiguous() and src.is_contiguous() and dst.numel() * dst.element_size() == src.numel() * src.element_size()
    src_ev = torch.cuda.Event()
    src_ev.record(torch.cuda.current_stream(src.device))
    with torch.cuda.device(dst.device):
        torch.cuda.current_stream(dst.device).wait_event(src_ev)
        check(require().pdb_peer_pull(ptr(dst), ptr(src), dst.numel() * dst.element_size(), _st(dst)), "pdb_peer_pull")
    count_launch()


def peer_copy(src: torch.Tensor, device: torch.device) -> torch.Tensor:
    dst = torch.empty(src.shape, dtype=src.dtype, device=device)
    peer_copy_(dst, src.contiguous())
    return dst


def reduce_to_root(root: torch.Tensor, peers) -> None:
    """root += sum(peers): GPU 0 reads every replica's gradient arena over NVLink (K4d)."""
    if not _has("pdb_reduce_to_root"):
        for g in peers:
            root.add_(g.to(root.device))
        return
    evs = []
    for g in peers:
        e = torch.cuda.Event()
        e.record(torch.cuda.current_stream(g.device))
        evs.append(e)
    with torch.cuda.device(root.device):
        for e in evs:
            torch.cuda.current_stream(root.device).wait_event(e)
        arr = (ctypes.c_void_p * len(peers))(*[g.data_ptr() for g in peers])
        check(require().pdb_reduce_to_root(ptr(root), arr, len(peers), root.numel(), _st(root)), "pdb_reduce_to_root")
    count_launch()


def peer_gather_(dst: torch.Tensor, srcs, shard_elems: int, skip: int = -1) -> None:
    """Shard q of flat `dst` <- the same range of srcs[q] (q != skip; None entries are skipped): the GPU of dst reads
    its peers over NVLink in ONE launch (second half of the sharded DataParallel replicate / reduce)."""
    es = dst.element_size()
    evs = []
    for i, t in enumerate(srcs):
        if t is None or i == skip:
            continue
        assert t.is_contiguous() and t.numel() == dst.numel() and t.dtype == dst.dtype
        e = torch.cuda.Event()
        e.record(torch.cuda.current_stream(t.device))
        evs.append(e)
    with torch.cuda.device(dst.device):
        cur = torch.cuda.current_stream(dst.device)
        for e in evs:
            cur.wait_event(e)
        arr = (ctypes.c_void_p * len(srcs))(*[0 if (t is None or i == skip) else t.data_ptr() for i, t in enumerate(srcs)])
        check(require().pdb_peer_gather(ptr(dst), arr, len(srcs), shard_elems * es, dst.numel() * es, skip, _st(dst)),
              "pdb_peer_gather")
    count_launch()
