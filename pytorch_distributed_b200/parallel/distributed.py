"""DistributedDataParallel over gradient arenas.

Constructor / attribute surface the reference uses (ref:restnet_ddp.py:99,
ref:resnet_ddp_apex.py:103): ``DistributedDataParallel(model.cuda(), device_ids=[local_rank])``
or without ``device_ids``; ``.module``; plus ``no_sync()`` and ``register_comm_hook`` from the
delegate (torch:nn/parallel/distributed.py:655-901,1474-1500,1987).

What is the same as the delegate
  * construction verifies parameter shapes across ranks and broadcasts rank 0's parameters
    and buffers (``_sync_module_states``, torch:distributed/utils.py:281-328);
  * every training forward re-broadcasts rank 0's BN buffers (``broadcast_buffers=True``,
    torch:nn/parallel/distributed.py:2214-2257);
  * gradients are averaged bucket by bucket, buckets are assigned walking parameters in reverse
    and launched strictly in order, overlapped with the rest of backward
    (reducer.hpp:276-336; 1 MiB first bucket, 25 MiB cap, reducer.hpp:30-31).

What is different
  * parameters and gradients live in arenas (utils/arena.py): a bucket is an address range, so
    there is no copy into a bucket before the collective and none out of it afterwards;
  * the bookkeeping runs in the C++ runtime (csrc/runtime/reducer.cpp);
  * with the native comm backend the collective is our peer-memory kernel, which also applies
    1/world, the amp un-scale and the inf check in the same pass (K1).
"""
from __future__ import annotations

import contextlib
import zlib
from typing import Optional

import torch
import torch.distributed as dist
from torch import nn
from torch.autograd import Variable

from .. import runtime
from ..utils.arena import ALIGN_ELEMS, BufferArena, ParamArena
from . import comm as _comm

_WIRE = {None: None, "fp32": None, "fp16": torch.float16, "bf16": torch.bfloat16,
         "allreduce": None, "fp16_compress": torch.float16, "bf16_compress": torch.bfloat16}


class GradBucket:
    """What a Python comm hook sees (subset of ``torch.distributed.GradBucket``): a view, never a copy."""

    def __init__(self, ddp: "DistributedDataParallel", index: int):
        self._ddp, self._index = ddp, index

    def index(self) -> int:
        return self._index

    def is_last(self) -> bool:
        return self._index == len(self._ddp.bucket_ranges) - 1

    def buffer(self) -> torch.Tensor:
        s, e = self._ddp.bucket_ranges[self._index]
        return self._ddp.arena.grads[s:e]

    def set_buffer(self, t: torch.Tensor) -> None:
        self.buffer().copy_(t.reshape(-1))

    def parameters(self):
        ps = list(self._ddp.arena.parameters())
        return [ps[i] for i, b in enumerate(self._ddp._param_bucket) if b == self._index]

    def gradients(self):
        return [p.grad for p in self.parameters()]


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, output_device=None, dim: int = 0,
                 broadcast_buffers: bool = True, process_group=None, bucket_cap_mb: float = 25.0,
                 first_bucket_mb: float = 1.0, find_unused_parameters: bool = False,
                 gradient_as_bucket_view: bool = True, backend: str = "auto",
                 comm_dtype: Optional[str] = None, init_sync: bool = True):
        super().__init__()
        if not _comm.is_initialized():
            raise RuntimeError("Default process group has not been initialized, "
                               "please make sure to call init_process_group.")
        if device_ids is not None and len(device_ids) > 1:
            raise ValueError("device_ids can only be None or contain a single element.")
        self.module = module
        self.device_ids = device_ids
        self.output_device = output_device if output_device is not None else (device_ids[0] if device_ids else None)
        self.dim = dim
        self.broadcast_buffers = broadcast_buffers
        self.process_group = process_group
        self.find_unused_parameters = find_unused_parameters
        self.require_backward_grad_sync = True
        self.require_forward_param_sync = True
        self.world_size = dist.get_world_size(process_group)
        self.rank = dist.get_rank(process_group)
        self.grad_scaler = None
        self._wire_dtype = _WIRE[comm_dtype]
        self._py_hook = None
        self._hook_results = []
        self._buf_event = None           # completion of an early (overlapped) buffer broadcast for the next forward

        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any "
                               "parameter that requires a gradient.")
        self.device = params[0].device
        self.arena: ParamArena = getattr(module, "_pdb200_arena", None) or ParamArena(module, self.device)
        self.buffer_arena: BufferArena = getattr(module, "_pdb200_buffer_arena", None) or BufferArena(module, self.device)
        self.comm = _comm.make_comm(backend, process_group)
        self.comm.register_arena(self.arena)

        if init_sync:
            self._verify_param_shapes()
            self._sync_module_states()

        offs = [e.offset for e in self.arena.entries]
        nums = [e.numel for e in self.arena.entries]
        self.bucket_ranges, self._param_bucket = runtime.plan_buckets(
            offs, nums, 4, int(first_bucket_mb * (1 << 20)), int(bucket_cap_mb * (1 << 20)), ALIGN_ELEMS)
        self.reducer = runtime.Reducer(self._param_bucket, len(self.bucket_ranges))
        self._pass_open = False
        self._param_index = {id(p): i for i, p in enumerate(self.arena.parameters())}
        self._hook_handles = []
        for i, p in enumerate(self.arena.parameters()):
            if p.requires_grad:
                self._hook_handles.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))
        eng = getattr(module, "_engine", None)
        if eng is not None:
            eng.attach_ddp(self)

    # -- construction-time collectives ---------------------------------------------------
    def _verify_param_shapes(self) -> None:
        desc = ";".join(f"{e.name}:{tuple(e.shape)}" for e in self.arena.entries).encode()
        h = torch.tensor([zlib.crc32(desc), len(self.arena.entries)], dtype=torch.int64, device=self._coll_device())
        ref = h.clone()
        dist.broadcast(ref, 0, group=self.process_group)
        ok = torch.tensor([int(torch.equal(h, ref))], dtype=torch.int64, device=h.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.process_group)
        if int(ok.item()) != 1:
            raise RuntimeError("DDP expects same model across all ranks, but parameter names/shapes differ "
                               f"on rank {self.rank}")

    def _coll_device(self) -> torch.device:
        return self.device if dist.get_backend(self.process_group) == "nccl" else torch.device("cpu")

    def _sync_module_states(self) -> None:
        self.comm.broadcast_(self.arena.params, 0)           # K2
        self.arena.refresh_shadow()
        self._sync_buffers()

    def _buffers_replicated(self) -> bool:
        """Native engine + SyncBatchNorm: every rank derives its running statistics from the same global sums, added
        in the same (rank) order, so the BN buffers are bit-identical everywhere after the construction-time broadcast
        and the per-iteration one (K3, torch:nn/parallel/distributed.py:2214-2257) would copy equal bytes."""
        eng = getattr(self.module, "_engine", None)
        return bool(eng is not None and getattr(eng, "sync_bn", False) and hasattr(self.comm, "sync_peers"))

    def _sync_buffers(self) -> None:                         # K3
        self.comm.broadcast_(self.buffer_arena.storage, 0)     # fp32 stats + int64 counters, one transfer

    # -- gradient path -----------------------------------------------------------------
    def _make_hook(self, index: int):
        def hook(param: torch.Tensor) -> None:
            # With the native step graph the engine reports readiness itself (mark_param_ready);
            # autograd still fires this hook for the graph's anchor parameter — ignore it.
            if getattr(self.module, "_engine", None) is None:
                self._on_grad_ready(index, param)
        return hook

    def _on_grad_ready(self, index: int, param=None) -> None:
        if not self.require_backward_grad_sync:
            return
        if param is not None and param.grad is not None:
            e = self.arena.entries[index]
            want = self.arena.grads.data_ptr() + 4 * e.offset
            if param.grad.data_ptr() != want:
                # A foreign zero_grad(set_to_none=True) dropped the alias: autograd allocated a fresh
                # gradient.  Move it into the arena and re-attach the view.
                view = self.arena.grad_view(e.name)
                view.copy_(param.grad)
                param.grad = view
        if not self._pass_open:
            self._pass_open = True
            self.reducer.reset()
            Variable._execution_engine.queue_callback(self._finalize_backward)
        for b in self.reducer.mark_ready(index):
            self._launch_bucket(b)

    def mark_param_ready(self, index: int) -> None:
        """Entry point for the native step graph, which knows statically when each gradient range
        has been written and does not go through per-parameter autograd hooks."""
        self._on_grad_ready(index, None)

    def _launch_bucket(self, b: int) -> None:
        s, e = self.bucket_ranges[b]
        if self._py_hook is not None:
            # user callback (torch:nn/parallel/distributed.py:1987): it owns the whole reduction of this range,
            # averaging included, and returns a tensor or a Future of one
            state, fn = self._py_hook
            bucket = GradBucket(self, b)
            self._hook_results.append((bucket, fn(state, bucket)))
            return
        sc = self.grad_scaler if (self.grad_scaler is not None and self.comm.fuses_unscale) else None
        self.comm.allreduce_range(
            self.arena.grads, s, e, average=True, wire_dtype=self._wire_dtype,
            inv_scale=sc.inv_scale_tensor(self.device) if sc is not None else None,
            found_inf=sc.found_inf_tensor(self.device) if sc is not None else None)

    def _finalize_backward(self) -> None:
        try:
            if not self.reducer.done():
                # Parameters that received no gradient keep a zero gradient in the arena; their buckets still
                # have to be reduced so that all ranks issue the same collectives.
                for b in self.reducer.flush():
                    self._launch_bucket(b)
            for bucket, res in self._hook_results:
                if not torch.is_tensor(res) and hasattr(res, "wait"):      # torch.futures.Future / torch._C.Future
                    res = res.wait()
                if isinstance(res, (list, tuple)):
                    res = res[0]
                buf = bucket.buffer()
                if res.data_ptr() != buf.data_ptr():
                    buf.copy_(res.reshape(-1))
            self._hook_results.clear()
            self.comm.wait_all()
            if self.grad_scaler is not None and self.comm.fuses_unscale and self._py_hook is None:
                self.grad_scaler.mark_unscaled_by_comm()
        finally:
            self._pass_open = False

    # -- public surface ------------------------------------------------------------------
    def attach_grad_scaler(self, scaler) -> None:
        """Let the fused all-reduce epilogue do the amp un-scale + inf check (K1/K13)."""
        self.grad_scaler = scaler

    def register_comm_hook(self, state, hook) -> None:
        """Built-in hooks — by name or by passing torch's ``allreduce_hook`` / ``fp16_compress_hook`` /
        ``bf16_compress_hook`` (default_comm_hooks.hpp:9-34, default_hooks.py:18-134) — are compile-time variants of
        the fused all-reduce kernel, not Python callbacks.  Any other callable is run as
        ``hook(state, bucket) -> Tensor | Future[Tensor]`` once per bucket with a :class:`GradBucket` whose
        ``buffer()`` is the bucket's range of the gradient arena; the result lands back in that range."""
        name = hook if isinstance(hook, str) else getattr(hook, "__name__", str(hook))
        name = name.replace("_hook", "")
        if name in _WIRE:
            self._wire_dtype, self._py_hook = _WIRE[name], None
        elif callable(hook):
            self._py_hook = (state, hook)
        else:
            raise ValueError(f"unsupported comm hook {name!r}; choose from allreduce, fp16_compress, bf16_compress "
                             "or pass a callable")

    @contextlib.contextmanager
    def no_sync(self):
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def forward(self, *inputs, **kwargs):
        with torch.autograd.profiler.record_function("pdb200.DistributedDataParallel.forward"):
            if torch.is_grad_enabled() and self.require_backward_grad_sync:
                self.require_forward_param_sync = True
            need_sync = self.require_forward_param_sync and self.broadcast_buffers and self.world_size > 1 \
                and not self._buffers_replicated()
            if self._buf_event is not None:
                # K3 was already run, on the comm stream, while the previous step's backward was executing (always
                # waited for, also when this forward would not have asked for it: the copy must not land mid-forward)
                torch.cuda.current_stream(self.device).wait_event(self._buf_event)
            elif need_sync:
                self._sync_buffers()
            self._buf_event = None
            if not torch.is_grad_enabled():
                # delegate clears the flag after a no_grad forward (distributed.py:1649-1650), so an
                # eval loop broadcasts buffers only on its first batch
                self.require_forward_param_sync = False
            if self.device_ids and inputs and torch.is_tensor(inputs[0]) and inputs[0].device != self.device:
                inputs = (inputs[0].to(self.device, non_blocking=True),) + tuple(inputs[1:])
            out = self.module(*inputs, **kwargs)
            if need_sync and torch.is_grad_enabled() and self.require_backward_grad_sync and hasattr(self.comm, "broadcast_async"):
                # The buffers only change in forward.  Rank 0's values — which the NEXT forward must start from
                # (broadcast_buffers=True, torch:nn/parallel/distributed.py:2214-2257) — are final once this forward
                # has run, so the broadcast is issued now, behind it, on the comm stream: it overlaps backward
                # instead of putting a cross-rank meeting at the top of every step.
                self._buf_event = self.comm.broadcast_async(self.buffer_arena.storage, 0)
            return out

    def load_state_dict(self, *args, **kwargs):
        self._buf_event = None                      # loaded buffers must be broadcast afresh at the next forward
        return super().load_state_dict(*args, **kwargs)

    def _get_ddp_logging_data(self):
        return {"bucket_sizes": ", ".join(str((e - s) * 4) for s, e in self.bucket_ranges),
                "num_buckets": len(self.bucket_ranges), "backend_name": self.comm.name,
                "world_size": self.world_size, "rank": self.rank}
