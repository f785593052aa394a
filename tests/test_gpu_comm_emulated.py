"""The peer-memory collectives with the ranks EMULATED on one GPU.

A multi-GPU box is not always at hand (and the driver's GPU test tier has one GPU), so every cross-rank kernel is also
run with `world` buffers, signal pads and staging areas that all live on cuda:0, one launch per emulated rank on a
stream of its own.  The kernels cannot tell the difference: they see `world` base pointers and meet on the pads with
system-scope release / acquire accesses exactly as they do over NVLink.  (Only the NVLS multicast variants need real
peers.)  Grids are kept small enough that the launches of all ranks are co-resident; where a rank's kernel fills the
GPU (the convolution), its non-meeting CTAs retire and make room for the peer, so the meeting still happens.
Oracle: plain torch sums of the per-rank inputs.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

from pytorch_distributed_b200 import ops  # noqa: E402
from pytorch_distributed_b200.ops import ConvShape, SyncPeersC  # noqa: E402
from pytorch_distributed_b200.ops._lib import check, ptr, require  # noqa: E402

from ._gpu_helpers import rel_err  # noqa: E402

PAD_WORDS = 64 * 16          # 64 block rows x up to 16 ranks


class Emu:
    """`world` emulated ranks: one buffer, one signal pad, one flag word and one stream each."""

    def __init__(self, world, n, seed=0, dtype=torch.float32):
        g = torch.Generator(device="cuda").manual_seed(seed)
        self.world = world
        self.bufs = [(torch.randn(n, device="cuda", generator=g) * (r + 1)).to(dtype) for r in range(world)]
        self.pads = [torch.zeros(PAD_WORDS, dtype=torch.int32, device="cuda") for _ in range(world)]
        self.flags = [torch.zeros(64, device="cuda") for _ in range(world)]
        self.streams = [torch.cuda.Stream() for _ in range(world)]
        torch.cuda.synchronize()

    @staticmethod
    def arr(ts):
        return (ctypes.c_void_p * len(ts))(*[t.data_ptr() for t in ts])

    def each(self, fn):
        """fn(rank, stream_ptr) once per rank, each on its own stream; returns after all have finished."""
        for r, st in enumerate(self.streams):
            with torch.cuda.stream(st):
                fn(r, ctypes.c_void_p(st.cuda_stream))
        torch.cuda.synchronize()


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("n,blocks", [(8 * 4096 + 64, 4), (1 << 20, 8)])
def test_two_shot_allreduce_emulated(world, n, blocks):
    L = require()
    e = Emu(world, n, seed=1)
    inv = torch.tensor(1.0 / 64.0, device="cuda")
    want = torch.stack(e.bufs).sum(0) / world / 64.0
    for rep in range(3):                       # repeated launches: the meeting counters keep counting
        if rep:
            for r in range(world):
                e.bufs[r].copy_(want * (r + 1))
            want = want * sum(range(1, world + 1)) / world / 64.0
        bufs, pads, flags = e.arr(e.bufs), e.arr(e.pads), e.arr(e.flags)
        e.each(lambda r, st: check(L.pdb_allreduce_inplace(bufs, pads, flags, None, r, world, 0, n, 1.0 / world, ptr(inv),
                                                           blocks, st), "pdb_allreduce_inplace"))
        for r in range(world):
            assert rel_err(e.bufs[r], want) < 1e-6, (rep, r)
            assert torch.equal(e.bufs[r], e.bufs[0])                  # replicas bit-identical
            assert float(e.flags[r][0]) == 0.0
    e.bufs[world - 1][777] = float("inf")                             # one rank overflows -> every rank must know
    e.each(lambda r, st: check(L.pdb_allreduce_inplace(bufs, pads, flags, None, r, world, 0, n, 1.0 / world, ptr(inv),
                                                       blocks, st), "pdb_allreduce_inplace"))
    assert all(float(f[0]) == 1.0 for f in e.flags)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("n", [64, 4096 + 4, 65536])
def test_one_shot_allreduce_emulated(world, n):
    L = require()
    e = Emu(world, n + 128, seed=2)
    inv = torch.tensor(0.5, device="cuda")
    start = 64
    want = torch.stack([b[start:start + n] for b in e.bufs]).sum(0) / world * 0.5
    before = [b.clone() for b in e.bufs]
    bufs, pads, flags = e.arr(e.bufs), e.arr(e.pads), e.arr(e.flags)
    e.each(lambda r, st: check(L.pdb_allreduce_oneshot(bufs, pads, flags, r, world, start, n, 1.0 / world, ptr(inv), st),
                               "pdb_allreduce_oneshot"))
    for r in range(world):
        assert rel_err(e.bufs[r][start:start + n], want) < 1e-6
        assert torch.equal(e.bufs[r][start:start + n], e.bufs[0][start:start + n])
        assert torch.equal(e.bufs[r][:start], before[r][:start]) and torch.equal(e.bufs[r][start + n:], before[r][start + n:])
        assert float(e.flags[r][0]) == 0.0
    e.bufs[0][start + 5] = float("nan")
    e.each(lambda r, st: check(L.pdb_allreduce_oneshot(bufs, pads, flags, r, world, start, n, 1.0 / world, ptr(inv), st),
                               "pdb_allreduce_oneshot"))
    assert all(float(f[0]) == 1.0 for f in e.flags)


@pytest.mark.parametrize("wire", [torch.bfloat16, torch.float16])
def test_wire16_allreduce_emulated(wire):
    L = require()
    world, n, blocks = 2, 1 << 18, 4
    e = Emu(world, n, seed=3)
    want = torch.stack(e.bufs).sum(0) / world
    half = n
    stage = [torch.zeros(2 * half, dtype=wire, device="cuda") for _ in range(world)]
    bufs, pads, flags, st16 = e.arr(e.bufs), e.arr(e.pads), e.arr(e.flags), e.arr(stage)
    e.each(lambda r, st: check(L.pdb_allreduce_inplace_wire16(bufs, pads, flags, st16, half, ops.dtype_tag(wire), r, world, 0, n,
                                                              1.0 / world, None, blocks, st), "pdb_allreduce_inplace_wire16"))
    for r in range(world):
        assert rel_err(e.bufs[r], want) < (2e-2 if wire == torch.bfloat16 else 2e-3)
        assert torch.equal(e.bufs[r], e.bufs[0])


def test_broadcast_emulated():
    L = require()
    world, n = 4, 1 << 16
    e = Emu(world, n, seed=4)
    src = e.bufs[2].clone()
    bufs, pads = e.arr(e.bufs), e.arr(e.pads)
    e.each(lambda r, st: check(L.pdb_broadcast_inplace(bufs, pads, r, world, 2, 0, n, 4, st), "pdb_broadcast_inplace"))
    assert all(torch.equal(b, src) for b in e.bufs)


def _sync_structs(world, slot, pads, staging):
    out = []
    for r in range(world):
        s = SyncPeersC()
        for q in range(world):
            s.staging[q] = staging[q].data_ptr()
            s.pads[q] = pads[q].data_ptr()
        s.rank, s.world, s.slot_stride, s.row = r, world, slot, 0
        out.append(s)
    return out


@pytest.mark.parametrize("world", [2, 8])
def test_small_allreduce_and_bn_stats_sync_emulated(world):
    L = require()
    slot = 1024
    e = Emu(world, 8, seed=5)
    staging = [torch.zeros(2 * world * slot, device="cuda") for _ in range(world)]
    stg, pads = e.arr(staging), e.arr(e.pads)
    for it in range(5):                                   # odd count: both halves of the staging area, reused
        vals = [torch.arange(516, device="cuda", dtype=torch.float32) * (r + 1) + it for r in range(world)]
        want = torch.stack(vals).sum(0)
        e.each(lambda r, st: check(L.pdb_small_allreduce(stg, pads, r, world, ptr(vals[r]), 516, slot, st), "pdb_small_allreduce"))
        assert all(torch.equal(v, want) for v in vals), it
    g = torch.Generator(device="cuda").manual_seed(6)
    C = 256
    xs = [(torch.randn(3000, C, device="cuda", generator=g) + r).to(torch.bfloat16) for r in range(world)]
    sums = [torch.empty(2 * C, device="cuda") for _ in range(world)]
    scratch = [torch.zeros(L.pdb_bn_stats_scratch_bytes(C), dtype=torch.uint8, device="cuda") for _ in range(world)]
    for _ in range(2):
        e.each(lambda r, st: check(L.pdb_bn_stats_sync(ptr(xs[r]), 0, 3000, C, ptr(sums[r]), ptr(scratch[r]), scratch[r].numel(),
                                                       stg, pads, r, world, slot, st), "pdb_bn_stats_sync"))
    want = sum(torch.cat([x.float().sum(0), (x.float() ** 2).sum(0)]) for x in xs)
    for r in range(world):
        assert rel_err(sums[r], want) < 1e-5 and torch.equal(sums[r], sums[0])


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float32])
def test_syncbn_exchange_inside_conv_and_bn_reduce_emulated(dt):
    """K6 on the hot path: the convolution's last CTA (forward) and the BN backward reduction's folding block exchange
    the [2C] sums themselves.  Two emulated ranks; oracle = statistics of the concatenated batch."""
    world, slot = 2, 4096
    N, H, C, K = 4, 28, 64, 256
    e = Emu(world, 8, seed=7)
    staging = [torch.zeros(2 * world * slot, device="cuda") for _ in range(world)]
    sync = _sync_structs(world, slot, e.pads, staging)
    g = torch.Generator(device="cuda").manual_seed(8)
    xs = [(torch.randn(N, H, H, C, device="cuda", generator=g) + 0.1 * r).to(dt) for r in range(world)]
    w = (torch.randn(K, 1, 1, C, device="cuda", generator=g) / C ** 0.5).to(dt)
    ys = [torch.empty(N, H, H, K, device="cuda", dtype=dt) for _ in range(world)]
    nscr = int(require().pdb_conv_stats_scratch_floats(K))
    for rep in range(3):
        stats = [torch.full((2 * K,), float("nan"), device="cuda") for _ in range(world)]
        scr = [torch.zeros(nscr, device="cuda") for _ in range(world)]

        def run(r, st):
            ops.conv_fprop(xs[r], w, ys[r], ConvShape(N, H, H, C, K, 1, 1, 1, 0), stats=stats[r], stats_scratch=scr[r], sync=sync[r])
        e.each(run)
        yall = torch.cat([y.float().reshape(-1, K) for y in ys])
        want = torch.cat([yall.sum(0), (yall * yall).sum(0)])
        for r in range(world):
            assert rel_err(stats[r], want) < 1e-4, (rep, r)
            assert torch.equal(stats[r], stats[0])                     # same numbers on every rank, bit for bit
    # backward: sum g and sum g*xhat over both ranks
    rows = N * H * H
    mean, rstd = torch.randn(K, device="cuda", generator=g) * 0.1, torch.rand(K, device="cuda", generator=g) + 0.5
    dz = [torch.randn(rows, K, device="cuda", generator=g).to(dt) for _ in range(world)]
    nb = ops.bn_bwd_reduce_scratch_floats(K)
    for rep in range(2):
        ds = [torch.full((2 * K,), float("nan"), device="cuda") for _ in range(world)]
        scr = [torch.zeros(nb, device="cuda") for _ in range(world)]

        def run_b(r, st):
            ops.bn_bwd_reduce(dz[r], None, ys[r].view(rows, K), mean, rstd, ds[r], scratch=scr[r], sync=sync[r])
        e.each(run_b)
        gall = torch.cat([d.float() for d in dz])
        xhat = (torch.cat([y.float().view(rows, K) for y in ys]) - mean) * rstd
        want = torch.cat([gall.sum(0), (gall * xhat).sum(0)])
        for r in range(world):
            assert rel_err(ds[r], want) < 2e-4 and torch.equal(ds[r], ds[0]), (rep, r)


def test_dp_peer_pull_and_reduce_to_root_one_device():
    """K4: the single-process DataParallel kernels (pull copy, reduce-to-root) — device-local pointers are valid
    arguments, which is all the kernels need."""
    g = torch.Generator(device="cuda").manual_seed(9)
    src = torch.randn(1 << 18, device="cuda", generator=g)
    dst = torch.empty_like(src)
    ops.peer_copy_(dst, src)
    torch.cuda.synchronize()
    assert torch.equal(dst, src)
    root = torch.randn(1 << 16, device="cuda", generator=g)
    peers = [torch.randn(1 << 16, device="cuda", generator=g) for _ in range(3)]
    want = root + sum(peers)
    ops.reduce_to_root(root, peers)
    torch.cuda.synchronize()
    assert rel_err(root, want) < 1e-6


def test_allreduce_trace_stamps_are_ordered():
    """--profile support: while a trace buffer is set, every two-shot launch stamps entry / first meeting / exit."""
    L = require()
    world, n = 2, 1 << 16
    e = Emu(world, n, seed=10)
    trace = torch.zeros(3 * 8, dtype=torch.int64, device="cuda")
    L.pdb_comm_set_trace(ptr(trace), 8)
    try:
        bufs, pads, flags = e.arr(e.bufs), e.arr(e.pads), e.arr(e.flags)
        for _ in range(2):
            e.each(lambda r, st: check(L.pdb_allreduce_inplace(bufs, pads, flags, None, r, world, 0, n, 1.0, None, 4, st), "allreduce"))
    finally:
        used = L.pdb_comm_set_trace(None, 0)
    assert used == 4
    t = trace.view(-1, 3).cpu()
    for i in range(4):
        a, b, c = (int(v) for v in t[i])
        assert 0 < a <= b <= c and c - a < 5_000_000_000, (i, a, b, c)
    assert int(t[4:].abs().sum()) == 0
