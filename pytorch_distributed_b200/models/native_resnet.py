"""Native sm_100a step graph for ResNet-50: the whole forward/backward as a fixed kernel sequence.

The reference's delegates execute ResNet-50 op by op through the dispatcher: ~160 forward and ~320
backward launches of cuDNN / ATen kernels per step, NCHW tensors, autograd bookkeeping per op
(SURVEY.md §3.2).  Here the network is a *static plan* over NHWC 16-bit buffers:

  conv  : tcgen05/TMA implicit GEMM (csrc/ops/conv_gemm.cu); its epilogue also produces the
          per-channel sum / sum-of-squares the following BatchNorm needs;
  stem  : the 7x7/2 convolution as a 4x4/1 convolution over the space-to-depth input, operands gathered by
          TMA im2col loads (no patch matrix; any other stem geometry falls back to an explicit im2col);
  BN    : one fused normalise(+residual)+ReLU pass forward, one reduce + one apply pass backward
          (csrc/ops/batchnorm.cu);
  dgrad : the same GEMM kernel on dY with flipped/transposed weights (stride-2 layers run on a
          zero-upsampled dY); wgrad: the MN-major split-K GEMM, accumulating straight into the fp32
          gradient arena (no per-parameter gradient tensors, nothing for DDP to flatten).

PyTorch sees ONE autograd node (``_NativeStep``).  Gradients are written into the ``ParamArena`` in
place; when the model is wrapped in our DistributedDataParallel the engine tells it, at statically
known points of the backward sequence, which parameters are complete so the bucket all-reduce can
start while earlier layers are still running — the role of the per-parameter autograd hooks in the
delegate's reducer (reducer.hpp:74,276-286), without the hooks.

Layer semantics follow tv:models/resnet.py:108-163,266-282 and torch:nn/modules/batchnorm.py:144-210.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch
from torch import nn

from .. import ops
from ..ops import ConvShape
from ..utils.arena import BufferArena, ParamArena

STEM_KPAD = 152          # 7*7*3 = 147 columns padded to a multiple of 8 (16-byte rows for TMA)


class _ConvPlan:
    """Static description of one convolution: where its weights live and its geometry."""

    def __init__(self, name: str, conv: nn.Conv2d, eng: "NativeResNetEngine"):
        self.name = name
        self.K, self.C = conv.out_channels, conv.in_channels
        self.R, self.S = conv.kernel_size
        self.stride, self.pad = conv.stride[0], conv.padding[0]
        self.pindex = eng.pindex[name + ".weight"]
        self.w16 = eng._weights_view(name + ".weight")          # [K, R, S, C] in the compute dtype (fp32: the master)
        self.dw = eng.arena.grads[eng.arena.by_name[name + ".weight"].offset:][: self.K * self.R * self.S * self.C]
        self.w_off = eng.arena.by_name[name + ".weight"].offset
        self.wt16 = None                      # slice of the engine's flat dgrad-weight buffer (set later)

    def shape(self, N: int, H: int, W: int) -> ConvShape:
        return ConvShape(N, H, W, self.C, self.K, self.R, self.S, self.stride, self.pad)

    def out_hw(self, H: int, W: int) -> Tuple[int, int]:
        return ((H + 2 * self.pad - self.R) // self.stride + 1, (W + 2 * self.pad - self.S) // self.stride + 1)


class _BNPlan:
    def __init__(self, name: str, bn: nn.BatchNorm2d, eng: "NativeResNetEngine", slot: int):
        self.name, self.C = name, bn.num_features
        self.eps = float(bn.eps)
        self.momentum = 0.1 if bn.momentum is None else float(bn.momentum)
        self.gamma = eng.arena.param_view(name + ".weight")
        self.beta = eng.arena.param_view(name + ".bias")
        self.dgamma = eng.arena.grad_view(name + ".weight")
        self.dbeta = eng.arena.grad_view(name + ".bias")
        self.pindex = (eng.pindex[name + ".weight"], eng.pindex[name + ".bias"])
        self.running_mean, self.running_var = bn.running_mean, bn.running_var
        self.nbt = bn.num_batches_tracked
        self.slot = slot                 # offset (floats) of this layer's [2C] block in the stats buffers
        self.rslot = eng._scratch_floats  # offset of this layer's replica scratch for the backward reduction
        self.rsize = ops.bn_bwd_reduce_scratch_floats(self.C)
        eng._scratch_floats += self.rsize
        self.mean = torch.empty(self.C, dtype=torch.float32, device=eng.device)
        self.rstd = torch.empty(self.C, dtype=torch.float32, device=eng.device)


class _BlockPlan:
    def __init__(self, prefix: str, block, eng: "NativeResNetEngine"):
        self.conv1, self.bn1 = eng._conv(prefix + ".conv1", block.conv1), eng._bn(prefix + ".bn1", block.bn1)
        self.conv2, self.bn2 = eng._conv(prefix + ".conv2", block.conv2), eng._bn(prefix + ".bn2", block.bn2)
        self.conv3, self.bn3 = eng._conv(prefix + ".conv3", block.conv3), eng._bn(prefix + ".bn3", block.bn3)
        self.convd = self.bnd = None
        if block.downsample is not None:
            self.convd = eng._conv(prefix + ".downsample.0", block.downsample[0])
            self.bnd = eng._bn(prefix + ".downsample.1", block.downsample[1])


class _NativeStep(torch.autograd.Function):
    """One autograd node for the whole network.  ``anchor`` is a parameter that requires grad; it
    only exists so that autograd schedules ``backward`` — real gradients go to the arena."""

    @staticmethod
    def forward(ctx, x, anchor, engine):
        logits, plan = engine._forward_train(x)
        ctx.engine, ctx.plan = engine, plan
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        ctx.engine._backward(ctx.plan, dlogits)
        ctx.plan = None
        return None, None, None


class _StepPlan:
    """Everything that is fixed for one input shape: the buffers of the step (allocated once, in a fixed
    order, then handed out again in the same order every step) and, once warm, the CUDA graphs of the forward
    pass and of each backward segment.  A training step then costs ~20 graph launches instead of ~350 kernel
    launches, which is what lets ONE Python process drive eight replicas in the DataParallel mode."""

    def __init__(self):
        self.bufs: List[torch.Tensor] = []
        self.cursor = 0
        self.steps = 0
        self.fwd_end = 0
        self.logits = None
        self.saved = None
        self.x_static = None
        self.dlogits = None
        self.fwd_graph = None
        self.bwd_graphs = None          # [(graph, ready parameter indices)]
        self.backward_ran = False
        self.graphs_failed = False


class NativeResNetEngine:
    def __init__(self, model, batch_size: int, dtype: torch.dtype = torch.bfloat16, sync_bn: bool = False,
                 deterministic: Optional[bool] = None):
        ops.require()
        p0 = next(model.parameters())
        if not p0.is_cuda:
            raise RuntimeError("the native engine needs the model on a CUDA device")
        if model.width % 64 != 0:
            raise RuntimeError(f"the tcgen05 conv path needs channel counts that are multiples of 64 (width={model.width})")
        if dtype not in (torch.bfloat16, torch.float16, torch.float32):
            raise RuntimeError("native engine compute dtype must be bf16, fp16 or fp32 (TF32 tensor-core math)")
        self.model, self.batch_size, self.dtype, self.sync_bn = model, batch_size, dtype, sync_bn
        self.device = p0.device
        self.arena: ParamArena = getattr(model, "_pdb200_arena", None) or ParamArena(model, self.device)
        self.buffers: BufferArena = getattr(model, "_pdb200_buffer_arena", None) or BufferArena(model, self.device)
        # GEMM operand source: the 16-bit shadow of the weights, or — fp32 storage / TF32 math — the master weights
        # themselves (already [Cout, kh, kw, Cin] in the arena; no copy, nothing to refresh after an optimizer step)
        if dtype == torch.float32:
            self.arena.shadow = None
            self._wsrc = self.arena.params
        else:
            self._wsrc = self.arena.ensure_shadow(dtype)
        self.pindex: Dict[str, int] = {e.name: i for i, e in enumerate(self.arena.entries)}
        self._stat_floats = 0
        self._scratch_floats = 0
        self.ddp = None
        import os
        # SyncBN: the per-layer statistics exchange is one of our peer-memory kernels (stateless flag protocol), so
        # it is captured and replayed like every other launch; with a torch.distributed fallback exchange the capture
        # aborts (see _sync_stats) and the step stays eager.
        self.graphs = os.environ.get("PDB200_GRAPHS", "1") != "0"
        # A failed capture is an error unless PDB200_GRAPHS_FALLBACK=1 asks for "warn and stay eager": a silent
        # fallback would quietly turn a 20-launch step into a 350-launch one.
        self.graphs_fallback = os.environ.get("PDB200_GRAPHS_FALLBACK", "0") == "1"
        self.graph_warmup = 2                     # eager steps per shape before capture
        # Deterministic mode (PDB200_DETERMINISTIC=1): every floating-point sum of the step is added in a fixed order —
        # BN statistics through the two-stage kernels instead of epilogue atomics, the BN backward reduction without
        # atomics, wgrad without split-K.  Slower (wgrad loses its parallelism over pixels), but two runs on the same
        # input give the same bits: "run twice, compare with torch.equal" is the race detector for the whole step,
        # and comparing against the default mode measures how much of a gradient is summation-order noise.
        self.deterministic = (os.environ.get("PDB200_DETERMINISTIC", "0") == "1") if deterministic is None else bool(deterministic)
        self._det_scratch = None
        self._splits = 1 if self.deterministic else 0          # wgrad split-K: 0 = heuristic, 1 = one CTA per tile
        self._plans: Dict[tuple, _StepPlan] = {}
        self._plan: Optional[_StepPlan] = None
        self._ready_log: Optional[List[int]] = None
        self._cap_stream = None

        self.stem_conv = self._conv("conv1", model.conv1)
        self.stem_bn = self._bn("bn1", model.bn1)
        self.stem_w16 = torch.zeros(model.conv1.out_channels, STEM_KPAD, dtype=dtype, device=self.device)
        # The standard 7x7/2 stem runs as a 4x4/1 convolution over the space-to-depth input, fed by TMA im2col loads
        # (no patch matrix); any other stem geometry goes through the explicit im2col matrix.
        c1 = self.stem_conv
        # (16-bit only: the tf32 instantiation of the stem goes through the explicit patch matrix)
        self.stem_s2d = (c1.R, c1.S, c1.stride, c1.pad, c1.C, c1.K) == (7, 7, 2, 3, 3, 64) and dtype != torch.float32
        self.stem_ws = torch.zeros(64, 256, dtype=dtype, device=self.device)
        self.stem_dws = torch.zeros(64, 256, dtype=torch.float32, device=self.device)
        self.blocks: List[_BlockPlan] = []
        for li in range(1, 5):
            layer = getattr(model, f"layer{li}")
            for bi, block in enumerate(layer):
                self.blocks.append(_BlockPlan(f"layer{li}.{bi}", block, self))
        fc = model.fc
        self.fc_K, self.fc_C = fc.out_features, fc.in_features
        self.fc_w16 = self._weights_view("fc.weight")
        self.fc_bias = self.arena.param_view("fc.bias")
        e = self.arena.by_name["fc.weight"]
        self.fc_dw = self.arena.grads[e.offset: e.offset + e.numel]
        self.fc_db = self.arena.grad_view("fc.bias")
        self.fc_pindex = (self.pindex["fc.weight"], self.pindex["fc.bias"])
        # forward: [sums of every BN | replica scratch + tickets for the conv epilogues], zeroed once per forward
        self.stats = torch.zeros(self._stat_floats + self._scratch_floats, dtype=torch.float32, device=self.device)
        # backward: [dsums of every BN | replica scratch + tickets of every BN], zeroed once per backward pass
        self.dstats = torch.zeros(self._stat_floats + self._scratch_floats, dtype=torch.float32, device=self.device)
        self._anchor = model.fc.bias
        if self.deterministic:
            maxc = max(b.C for b in [self.stem_bn] + [x for blk in self.blocks for x in (blk.bn1, blk.bn2, blk.bn3, blk.bnd) if x is not None])
            self._det_scratch = torch.zeros(ops.require().pdb_bn_stats_scratch_bytes(maxc), dtype=torch.uint8, device=self.device)
        self._convs: List[_ConvPlan] = [self.stem_conv] + [c for b in self.blocks for c in
                                                           (b.conv1, b.conv2, b.conv3, b.convd) if c is not None]
        # dgrad operand layouts [Cin, kh, kw(flipped), Cout] of every conv but the stem, plus fc, in ONE flat
        # buffer so a single batched launch refreshes them all from the weight shadow
        table, off = [], 0
        for cv in self._convs[1:]:
            n = cv.K * cv.R * cv.S * cv.C
            table.append([cv.w_off, off, cv.K, cv.R, cv.S, cv.C])
            off += (n + 63) // 64 * 64
        fc_off = off
        table.append([e.offset, fc_off, self.fc_K, 1, 1, self.fc_C])
        off += (self.fc_K * self.fc_C + 63) // 64 * 64
        self.wt_all = torch.empty(off, dtype=dtype, device=self.device)
        for cv, row in zip(self._convs[1:], table):
            cv.wt16 = self.wt_all[row[1]: row[1] + cv.K * cv.R * cv.S * cv.C]
        self.fc_wt16 = self.wt_all[fc_off: fc_off + self.fc_K * self.fc_C]
        self._wt_table = torch.tensor(table, dtype=torch.int64, device=self.device)

    # -- plan construction -------------------------------------------------------------------------
    def _weights_view(self, name: str) -> torch.Tensor:
        e = self.arena.by_name[name]
        return self._wsrc[e.offset: e.offset + e.numel]

    def _conv(self, name: str, conv: nn.Conv2d) -> _ConvPlan:
        return _ConvPlan(name, conv, self)

    def _bn(self, name: str, bn: nn.BatchNorm2d) -> _BNPlan:
        plan = _BNPlan(name, bn, self, self._stat_floats)
        self._stat_floats += 2 * bn.num_features
        return plan

    def attach_ddp(self, ddp) -> None:
        self.ddp = ddp
        if self.sync_bn and getattr(ddp, "world_size", 1) > 1 and not hasattr(getattr(ddp, "comm", None), "allreduce_small_"):
            # SyncBN statistics would go through torch.distributed calls (explicit --backend nccl): those are not
            # capturable, so this configuration runs the step eagerly — decided here, not discovered by a failed capture
            self.graphs = False

    def _ready(self, *indices: int) -> None:
        if self._ready_log is not None:           # capturing: remember, tell DDP when the segment is replayed
            self._ready_log.extend(indices)
        elif self.ddp is not None:
            for i in indices:
                self.ddp.mark_param_ready(i)

    # -- helpers -------------------------------------------------------------------------------------
    def _new(self, *shape, dtype: Optional[torch.dtype] = None) -> torch.Tensor:
        """Step buffers come from the current plan: allocated on the first step of a shape, re-used (same
        order, same addresses) afterwards — no allocator traffic in the step and capturable in a CUDA graph."""
        dtype = dtype or self.dtype
        p = self._plan
        if p is None:
            return torch.empty(shape, dtype=dtype, device=self.device)
        if p.cursor < len(p.bufs):
            t = p.bufs[p.cursor]
            if tuple(t.shape) != tuple(shape) or t.dtype != dtype:
                raise RuntimeError("native engine: the step's buffer sequence changed between steps "
                                   f"(slot {p.cursor}: {tuple(t.shape)} vs {tuple(shape)})")
        else:
            t = torch.empty(shape, dtype=dtype, device=self.device)
            p.bufs.append(t)
        p.cursor += 1
        return t

    def _capture(self, fn):
        """Captures fn's launches into a CUDA graph; the returned graph remembers how many of our kernels it
        holds so that replays keep the library's launch counter honest."""
        from ..ops import _lib
        n0 = _lib.thread_launch_count()
        g = torch.cuda.CUDAGraph()
        if self._cap_stream is None:
            self._cap_stream = torch.cuda.Stream(self.device)
        cur = torch.cuda.current_stream(self.device)
        self._cap_stream.wait_stream(cur)
        with torch.cuda.graph(g, stream=self._cap_stream, capture_error_mode="thread_local"):
            fn()
        cur.wait_stream(self._cap_stream)
        g.pdb200_kernels = _lib.thread_launch_count() - n0
        return g

    @staticmethod
    def _replay(g) -> None:
        from ..ops import _lib
        g.replay()
        _lib.count_launch(getattr(g, "pdb200_kernels", 0))

    def _refresh_weights(self, need_dgrad: bool) -> None:
        """16-bit operand copies derived from the weight shadow: padded stem rows and, for training,
        the flipped/transposed layouts dgrad runs on."""
        c = self.stem_conv
        if self.stem_s2d:
            ops.stem_weight_s2d(c.w16, self.stem_ws)
        ops.pad_rows16(c.w16, self.stem_w16, c.K, c.R * c.S * c.C, STEM_KPAD)
        if need_dgrad:
            ops.weight_dgrad_layout_batched(self._wsrc, self.wt_all, self._wt_table, self._wt_table.shape[0])

    def _fwd_scratch(self, bn: _BNPlan) -> torch.Tensor:
        r0 = self._stat_floats + bn.rslot
        return self.stats[r0: r0 + bn.rsize]

    def _sync_world(self) -> int:
        if not self.sync_bn:
            return 1
        import torch.distributed as dist
        return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1

    def _sync_handle(self, n_floats: int):
        """SyncBatchNorm on the native comm backend: the handle that lets the kernel which already holds the [2C]
        sums (conv epilogue forward, BN reduction backward) add them up over the GPUs itself — K6, zero extra
        launches.  None: no SyncBN, one rank, or a backend without peer memory (then `_sync_stats` does it)."""
        if self._sync_world() <= 1:
            return None
        comm = getattr(self.ddp, "comm", None)
        return comm.sync_peers(n_floats) if hasattr(comm, "sync_peers") else None

    def _sync_stats(self, buf: torch.Tensor, fused: bool) -> float:
        """SyncBatchNorm: sums are additive, so one all-reduce of [2C] per layer and direction.  `fused`: the producing
        kernel has already exchanged them.  Returns the factor on the per-channel element count."""
        world = self._sync_world()
        if world <= 1:
            return 1.0
        if not fused:
            import torch.distributed as dist
            comm = getattr(self.ddp, "comm", None)
            if comm is not None and hasattr(comm, "allreduce_small_"):
                comm.allreduce_small_(buf)        # peer-memory exchange as a launch of its own (payload > one slot)
            else:
                if buf.is_cuda and torch.cuda.is_current_stream_capturing():
                    raise RuntimeError("SyncBN statistics go through torch.distributed here: not captured")
                dist.all_reduce(buf)
        return float(world)

    def _conv_bn(self, cv: _ConvPlan, bn: _BNPlan, x: torch.Tensor, *, residual=None, relu=True, train=True):
        N, H, W, _ = x.shape
        P, Q = cv.out_hw(H, W)
        y = self._new(N, P, Q, cv.K)
        sums = self.stats[bn.slot: bn.slot + 2 * bn.C] if train else None
        sync = self._sync_handle(2 * bn.C) if train and not self.deterministic else None
        fused_stats = train and not self.deterministic
        ops.conv_fprop(x, cv.w16, y, cv.shape(N, H, W), stats=sums if fused_stats else None,
                       stats_scratch=self._fwd_scratch(bn) if fused_stats else None, sync=sync)
        z = self._new(N, P, Q, cv.K)
        rows = N * P * Q
        if train:
            if not fused_stats:
                ops.bn_stats(y.view(rows, cv.K), sums, self._det_scratch)       # fixed summation order
            world = self._sync_stats(sums, fused=sync is not None)
            # one bit per element records the ReLU decision; BN backward reads it instead of z (1/16 the bytes)
            mask = self._new(rows * cv.K // 8, dtype=torch.uint8) if relu else None
            ops.bn_apply(y.view(rows, cv.K), None if residual is None else residual.view(rows, cv.K),
                         z.view(rows, cv.K), sums, rows * world, bn.gamma, bn.beta, bn.eps, bn.momentum,
                         bn.running_mean, bn.running_var, bn.nbt, bn.mean, bn.rstd, relu, mask)
            return y, z, mask
        else:
            ops.bn_apply_eval(y.view(rows, cv.K), None if residual is None else residual.view(rows, cv.K),
                              z.view(rows, cv.K), bn.running_mean, bn.running_var, bn.gamma, bn.beta, bn.eps, relu)
        return y, z, None

    # -- forward -----------------------------------------------------------------------------------------
    def _stem(self, x: torch.Tensor, train: bool):
        N, _, H, W = x.shape
        cv, bn = self.stem_conv, self.stem_bn
        P, Q = cv.out_hw(H, W)
        rows = N * P * Q
        y = self._new(N, P, Q, cv.K)
        sums = self.stats[bn.slot: bn.slot + 2 * bn.C] if train else None
        sync = self._sync_handle(2 * bn.C) if train and not self.deterministic else None
        fused_stats = train and not self.deterministic
        if not fused_stats:
            sums_arg, scr_arg = None, None
        else:
            sums_arg, scr_arg = sums, self._fwd_scratch(bn)
        if self.stem_s2d and H % 2 == 0 and W % 2 == 0:
            col = self._new(N, H // 2, W // 2, 16)          # space-to-depth input; also the wgrad operand
            ops.stem_s2d(x.contiguous().float(), col)
            ops.stem_fprop(col, self.stem_ws, y, stats=sums_arg, stats_scratch=scr_arg, sync=sync)
        else:
            col = self._new(rows, STEM_KPAD)
            ops.stem_im2col(x.contiguous().float(), col)
            ops.conv_fprop(col, self.stem_w16, y, ConvShape(rows, 1, 1, STEM_KPAD, cv.K, 1, 1, 1, 0), stats=sums_arg,
                           stats_scratch=scr_arg, sync=sync)
        z = self._new(N, P, Q, cv.K)
        mask = None
        if train:
            if not fused_stats:
                ops.bn_stats(y.view(rows, cv.K), sums, self._det_scratch)
            world = self._sync_stats(sums, fused=sync is not None)
            mask = self._new(rows * cv.K // 8, dtype=torch.uint8)
            ops.bn_apply(y.view(rows, cv.K), None, z.view(rows, cv.K), sums, rows * world, bn.gamma, bn.beta, bn.eps,
                         bn.momentum, bn.running_mean, bn.running_var, bn.nbt, bn.mean, bn.rstd, True, mask)
        else:
            ops.bn_apply_eval(y.view(rows, cv.K), None, z.view(rows, cv.K), bn.running_mean, bn.running_var, bn.gamma,
                              bn.beta, bn.eps, True)
        Hp, Wp = (P + 2 - 3) // 2 + 1, (Q + 2 - 3) // 2 + 1
        pooled = self._new(N, Hp, Wp, cv.K)
        argmax = self._new(N, Hp, Wp, cv.K, dtype=torch.uint8)
        ops.maxpool_fwd(z, pooled, argmax)
        return col, y, z, argmax, mask, pooled

    def _block_fwd(self, b: _BlockPlan, x: torch.Tensor, train: bool):
        y1, z1, m1 = self._conv_bn(b.conv1, b.bn1, x, train=train)
        y2, z2, m2 = self._conv_bn(b.conv2, b.bn2, z1, train=train)
        if b.convd is not None:
            yd, sc, _ = self._conv_bn(b.convd, b.bnd, x, relu=False, train=train)
        else:
            yd, sc = None, x
        y3, out, m3 = self._conv_bn(b.conv3, b.bn3, z2, residual=sc, relu=True, train=train)
        return (x, y1, z1, m1, y2, z2, m2, y3, yd, m3), out

    def _head(self, feat: torch.Tensor):
        N, H, W, C = feat.shape
        pooled = self._new(N, C)
        ops.avgpool_fwd(feat, pooled)
        logits = self._new(N, self.fc_K)
        ops.conv_fprop(pooled, self.fc_w16, logits, ConvShape(N, 1, 1, C, self.fc_K, 1, 1, 1, 0), bias=self.fc_bias,
                       block_n=128)
        return pooled, logits

    def _forward_impl(self, x: torch.Tensor):
        self.stats.zero_()
        self._refresh_weights(need_dgrad=True)
        stem = self._stem(x, True)
        cur, saved_blocks = stem[-1], []
        for b in self.blocks:
            sv, cur = self._block_fwd(b, cur, True)
            saved_blocks.append(sv)
        pooled, logits = self._head(cur)
        return logits, (stem, saved_blocks, pooled, tuple(cur.shape))

    def _forward_train(self, x: torch.Tensor):
        x = x.contiguous()
        if x.dtype != torch.float32:
            x = x.float()
        key = (tuple(x.shape), "train")
        plan = self._plans.get(key)
        if plan is None:
            plan = self._plans[key] = _StepPlan()
        self._plan = plan
        plan.cursor = 0
        plan.steps += 1
        try:
            if plan.fwd_graph is None and self.graphs and not plan.graphs_failed and plan.backward_ran \
                    and plan.steps > self.graph_warmup:
                try:
                    plan.x_static = x.clone()

                    def fwd():
                        plan.cursor = 0
                        plan.logits, plan.saved = self._forward_impl(plan.x_static)
                        plan.fwd_end = plan.cursor
                    plan.fwd_graph = self._capture(fwd)
                except Exception as e:
                    if not self.graphs_fallback:
                        raise RuntimeError("pdb200: CUDA-graph capture of the forward pass failed (set PDB200_GRAPHS=0 "
                                           f"to run eagerly, PDB200_GRAPHS_FALLBACK=1 to fall back): {e}") from e
                    import warnings
                    warnings.warn(f"pdb200: CUDA-graph capture of the forward pass failed, staying eager: {e}")
                    plan.fwd_graph, plan.graphs_failed = None, True
                    plan.cursor = 0               # the aborted capture consumed buffer slots: hand them out again
            if plan.fwd_graph is not None:
                plan.x_static.copy_(x, non_blocking=True)
                self._replay(plan.fwd_graph)
                plan.cursor = plan.fwd_end
            else:
                plan.logits, plan.saved = self._forward_impl(x)
                plan.fwd_end = plan.cursor
        finally:
            self._plan = None
        # autograd needs a tensor of its own per step (the static logits buffer is overwritten by the next one)
        return plan.logits.clone(), plan

    @torch.no_grad()
    def forward_eval(self, x: torch.Tensor, train_stats: bool = False) -> torch.Tensor:
        if train_stats:
            self.stats.zero_()
        self._refresh_weights(need_dgrad=False)
        cur = self._stem(x, train_stats)[-1]
        for b in self.blocks:
            _, cur = self._block_fwd(b, cur, train_stats)
        return self._head(cur)[1]

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        if self.model.training and torch.is_grad_enabled():
            return _NativeStep.apply(x, self._anchor, self)
        return self.forward_eval(x, train_stats=self.model.training)

    # -- backward ----------------------------------------------------------------------------------------
    def _bn_bwd(self, bn: _BNPlan, dz, mask, y, *, want_dres: bool):
        """Returns (dy, dres).  dres = dz * relu_mask (gradient of the residual branch); `mask` is the bit mask
        written by the forward pass (None for a BN that is not followed by a ReLU)."""
        rows = y.numel() // bn.C
        ds = self.dstats[bn.slot: bn.slot + 2 * bn.C]
        v = lambda t: None if t is None else t.view(rows, bn.C)    # noqa: E731
        r0 = self._stat_floats + bn.rslot
        scratch = self.dstats[r0: r0 + bn.rsize]
        if self.deterministic:
            sync = None
            ops.bn_bwd_reduce_det(v(dz), None, v(y), bn.mean, bn.rstd, ds, relu_mask=mask, scratch=self._det_scratch)
        else:
            sync = self._sync_handle(2 * bn.C)
            ops.bn_bwd_reduce(v(dz), None, v(y), bn.mean, bn.rstd, ds, relu_mask=mask, scratch=scratch, sync=sync)
        world = self._sync_stats(ds, fused=sync is not None)
        dy = self._new(*y.shape)
        dres = self._new(*y.shape) if want_dres else None
        ops.bn_bwd_apply(v(dz), None, v(y), v(dy), v(dres), bn.mean, bn.rstd, bn.gamma, ds, rows * world, bn.dgamma,
                         bn.dbeta, relu_mask=mask)
        return dy, dres

    def _conv_bwd(self, cv: _ConvPlan, x: torch.Tensor, dy: torch.Tensor, need_dx: bool = True, into=None):
        """wgrad into the arena; returns dx (same spatial size as x) when requested.  `into`: an existing
        gradient buffer of dx's shape that the dgrad result is ADDED to in the GEMM epilogue (TMA reduction
        store) — how the two branches of a residual block are summed without a separate elementwise pass."""
        N, H, W, _ = x.shape
        ops.conv_wgrad(x, dy, cv.dw, cv.shape(N, H, W), splits=self._splits)
        if not need_dx:
            return None
        _, P, Q, _ = dy.shape
        if cv.R == 1:
            # 1x1: plain GEMM on the (possibly strided-compact) pixel set
            dx = into if into is not None else self._new(N, P, Q, cv.C)
            ops.conv_fprop(dy, cv.wt16, dx, ConvShape(N, P, Q, cv.K, cv.C, 1, 1, 1, 0), accumulate=into is not None)
            return dx            # compact [N, P, Q, C]; the caller scatters it if stride == 2
        if cv.stride == 2:
            up = self._new(N, H, W, cv.K)
            ops.upsample2_zero(dy, up)
            dy, P, Q = up, H, W
        dx = self._new(N, H, W, cv.C)
        ops.conv_fprop(dy, cv.wt16, dx, ConvShape(N, P, Q, cv.K, cv.C, cv.R, cv.S, 1, cv.R - 1 - cv.pad))
        return dx

    def _block_bwd(self, b: _BlockPlan, sv, dout: torch.Tensor) -> torch.Tensor:
        x, y1, z1, m1, y2, z2, m2, y3, yd, m3 = sv
        dy3, dres = self._bn_bwd(b.bn3, dout, m3, y3, want_dres=True)
        self._ready(*b.bn3.pindex)
        dz2 = self._conv_bwd(b.conv3, z2, dy3)
        self._ready(b.conv3.pindex)
        dy2, _ = self._bn_bwd(b.bn2, dz2, m2, y2, want_dres=False)
        self._ready(*b.bn2.pindex)
        dz1 = self._conv_bwd(b.conv2, z1, dy2)
        self._ready(b.conv2.pindex)
        dy1, _ = self._bn_bwd(b.bn1, dz1, m1, y1, want_dres=False)
        self._ready(*b.bn1.pindex)
        if b.convd is None:
            # identity shortcut: dx = dgrad(conv1) + dres, summed in the GEMM epilogue on top of dres
            dx = self._conv_bwd(b.conv1, x, dy1, into=dres)
            self._ready(b.conv1.pindex)
            return dx
        dxa = self._conv_bwd(b.conv1, x, dy1)
        self._ready(b.conv1.pindex)
        dyd, _ = self._bn_bwd(b.bnd, dres, None, yd, want_dres=False)
        self._ready(*b.bnd.pindex)
        if b.convd.stride == 2:
            dxd = self._conv_bwd(b.convd, x, dyd)
            self._ready(b.convd.pindex)
            dx = self._new(*dxa.shape)
            ops.add_strided2(dxa, dxd, dx)
            return dx
        dx = self._conv_bwd(b.convd, x, dyd, into=dxa)       # stride-1 projection (layer1.0): same trick
        self._ready(b.convd.pindex)
        return dx

    # The backward pass as a list of segments (head, one per bottleneck in reverse, stem).  Each segment is a
    # CUDA graph once warm; DDP is told about a segment's parameters right after the segment has been enqueued,
    # so bucket all-reduces still overlap with the remaining segments.
    def _seg_head(self, plan: _StepPlan, st: dict) -> None:
        _, _, pooled, feat_shape = plan.saved
        dl = plan.dlogits
        N = dl.shape[0]
        self.dstats.zero_()
        ops.conv_wgrad(pooled, dl, self.fc_dw, ConvShape(N, 1, 1, self.fc_C, self.fc_K, 1, 1, 1, 0), splits=self._splits)
        ops.colsum16(dl, self.fc_db)
        self._ready(*self.fc_pindex[::-1])
        dpooled = self._new(N, self.fc_C)
        ops.conv_fprop(dl, self.fc_wt16, dpooled, ConvShape(N, 1, 1, self.fc_K, self.fc_C, 1, 1, 1, 0))
        d = self._new(*feat_shape)
        ops.avgpool_bwd(dpooled, d)
        st["d"] = d

    def _seg_block(self, i: int):
        def run(plan: _StepPlan, st: dict) -> None:
            st["d"] = self._block_bwd(self.blocks[i], plan.saved[1][i], st["d"])
        return run

    def _seg_stem(self, plan: _StepPlan, st: dict) -> None:
        col, y0, z0, argmax, m0, _ = plan.saved[0]
        dz0 = self._new(*z0.shape)
        ops.maxpool_bwd(st["d"], argmax, dz0)
        dy0, _ = self._bn_bwd(self.stem_bn, dz0, m0, y0, want_dres=False)
        self._ready(*self.stem_bn.pindex)
        cv = self.stem_conv
        if col.dim() == 4:                                  # space-to-depth path
            self.stem_dws.zero_()
            ops.stem_wgrad(col, dy0, self.stem_dws, splits=self._splits)
            ops.stem_dw_unpack(self.stem_dws, cv.dw, accumulate=True)
        else:
            rows = col.shape[0]
            ops.conv_wgrad(col, dy0.view(rows, cv.K), cv.dw, ConvShape(rows, 1, 1, STEM_KPAD, cv.K, 1, 1, 1, 0),
                           ldw=cv.R * cv.S * cv.C, kvalid=cv.R * cv.S * cv.C, splits=self._splits)
        self._ready(cv.pindex)

    def _segments(self):
        return [self._seg_head] + [self._seg_block(i) for i in reversed(range(len(self.blocks)))] + [self._seg_stem]

    def _backward(self, plan: _StepPlan, dlogits: torch.Tensor) -> None:
        self._plan = plan
        try:
            plan.cursor = plan.fwd_end
            plan.dlogits = self._new(*dlogits.shape)
            plan.dlogits.copy_(dlogits)                   # also converts an fp32 incoming gradient
            if plan.bwd_graphs is None and plan.fwd_graph is not None and not plan.graphs_failed:
                try:
                    graphs, st = [], {}
                    for seg in self._segments():
                        self._ready_log = []
                        g = self._capture(lambda seg=seg: seg(plan, st))
                        graphs.append((g, tuple(self._ready_log)))
                    plan.bwd_graphs = graphs
                except Exception as e:
                    if not self.graphs_fallback:
                        raise RuntimeError("pdb200: CUDA-graph capture of the backward pass failed (set PDB200_GRAPHS=0 "
                                           f"to run eagerly, PDB200_GRAPHS_FALLBACK=1 to fall back): {e}") from e
                    import warnings
                    warnings.warn(f"pdb200: CUDA-graph capture of the backward pass failed, staying eager: {e}")
                    plan.bwd_graphs, plan.graphs_failed = None, True
                finally:
                    self._ready_log = None
                    plan.cursor = plan.fwd_end + 1
            if plan.bwd_graphs is not None:
                for g, ready in plan.bwd_graphs:
                    self._replay(g)
                    self._ready(*ready)
            else:
                st: dict = {}
                for seg in self._segments():
                    seg(plan, st)
            plan.backward_ran = True
        finally:
            self._plan = None
