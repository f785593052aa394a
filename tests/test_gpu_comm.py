"""Peer-memory collectives and the DDP / DP engines on >= 2 GPUs, against NCCL as the oracle."""
import json

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]

from pytorch_distributed_b200 import launcher, ops  # noqa: E402

from . import _mp_gpu as W  # noqa: E402


def _world():
    return min(torch.cuda.device_count(), 8)


@pytest.mark.timeout(600)
def test_collectives_match_nccl(tmp_path, free_port):
    world = _world()
    launcher.spawn(W.comm_worker, args=(world, free_port, str(tmp_path)), nprocs=world)
    for r in range(world):
        res = json.load(open(tmp_path / f"comm_{r}.json"))
        for k, v in res.items():
            if k.startswith("allreduce_err"):
                assert v < 1e-5, (k, v, res)          # fp32 sums in a fixed order: only association differs
            elif k.startswith("found_inf_clean"):
                assert v == 0.0, res
            elif k.startswith("found_inf_injected"):
                assert v == 1.0, res                  # ... on every rank, not only the one that saw the inf
        assert res["broadcast_ok"] and res["small_allreduce_ok"] and res["soak_ok"], res
        assert res["wire_bf16_err"] < 2e-2 and res["wire_fp16_err"] < 2e-3, res      # 16-bit wire rounding only
        assert res["bn_stats_sync_err"] < 1e-4, res


@pytest.mark.timeout(900)
@pytest.mark.parametrize("sync_bn", [False, True])
def test_ddp_native_comm_matches_nccl_backend(tmp_path, free_port, sync_bn):
    world = _world()
    for i, backend in enumerate(("native", "nccl")):
        launcher.spawn(W.ddp_native_worker, args=(world, free_port + i, str(tmp_path), backend, sync_bn), nprocs=world)
    nat = [torch.load(tmp_path / f"ddp_native_{int(sync_bn)}_{r}.pt") for r in range(world)]
    ref = [torch.load(tmp_path / f"ddp_nccl_{int(sync_bn)}_{r}.pt") for r in range(world)]
    assert nat[0]["comm"] == "native" and ref[0]["comm"] == "torch"
    # --profile bookkeeping: one timed entry per bucket and step, with a sane bandwidth figure
    bt = nat[0]["buckets"]
    assert bt and len(bt) % 5 == 0 and all(b["us"] > 0 and b["bus_gbs"] > 0 and b["bytes"] > 0 for b in bt)
    for r in range(1, world):
        assert torch.equal(nat[0]["params"], nat[r]["params"])        # replicas stay bit-identical
    for r in range(world):
        # 16-bit rounding makes step >= 1 mildly chaotic at this tiny batch; the replicas-identical check above
        # is the exact one
        assert nat[r]["losses"][:3] == pytest.approx(ref[r]["losses"][:3], rel=6e-2)
        assert nat[r]["losses"] == pytest.approx(ref[r]["losses"], rel=0.2)      # steps 3, 4 replay the CUDA graphs
        assert nat[r]["scale"] == 65536.0
    cos = torch.nn.functional.cosine_similarity(nat[0]["params"] - ref[0]["params"].mean() * 0, ref[0]["params"], dim=0)
    assert float(cos) > 0.9999


@pytest.mark.timeout(900)
def test_ddp_native_comm_matches_nccl_backend_fp32_deterministic(tmp_path, free_port):
    """The `ddp` mode at the reference's precision on the deterministic engine: native peer-memory all-reduce vs NCCL
    differ only in how the cross-rank sum is associated — five steps (two eager, three through CUDA graphs) agree to 1e-5
    in every loss (measured: 1e-7); replicas stay bit-identical."""
    world = _world()
    for i, backend in enumerate(("native", "nccl")):
        launcher.spawn(W.ddp_native_worker, args=(world, free_port + i, str(tmp_path), backend, False, True), nprocs=world)
    nat = [torch.load(tmp_path / f"ddp_native_0_precise_{r}.pt") for r in range(world)]
    ref = [torch.load(tmp_path / f"ddp_nccl_0_precise_{r}.pt") for r in range(world)]
    assert nat[0]["comm"] == "native" and ref[0]["comm"] == "torch"
    for r in range(1, world):
        assert torch.equal(nat[0]["params"], nat[r]["params"])
    for r in range(world):
        print(f"rank {r} losses native {nat[r]['losses']} nccl {ref[r]['losses']}")
        assert nat[r]["losses"] == pytest.approx(ref[r]["losses"], rel=1e-5)
    d = (nat[0]["params"] - ref[0]["params"]).norm() / ref[0]["params"].norm()
    assert float(d) < 1e-4, float(d)


def test_dp_peer_kernels():
    d0, d1 = torch.device("cuda", 0), torch.device("cuda", 1)
    src = torch.randn(1 << 20, device=d0)
    dst = ops.peer_copy(src, d1)
    torch.cuda.synchronize(d1)
    assert torch.equal(dst.cpu(), src.cpu())
    root = torch.randn(1 << 20, device=d0)
    peers = [torch.randn(1 << 20, device=torch.device("cuda", i)) for i in range(1, torch.cuda.device_count())]
    want = root.cpu() + sum(p.cpu() for p in peers)
    ops.reduce_to_root(root, peers)
    torch.cuda.synchronize(d0)
    assert torch.allclose(root.cpu(), want, rtol=1e-5, atol=1e-5)


@pytest.mark.timeout(600)
@pytest.mark.parametrize("dt,det,shard_min", [(torch.float32, True, "3"), (torch.float32, True, "2"), (torch.float16, False, "3")])
def test_data_parallel_native_two_gpus(dt, det, shard_min, monkeypatch):
    """DP mode with the native engine per replica: gradients are the SUM over replicas on GPU 0
    (ref:resnet_dp.py:24,82).  In deterministic fp32 mode every sum has a fixed order, so the two-replica result must
    equal the sequential oracle (same two chunks through one engine) to rounding of a single addition; the fp16
    default-mode case is held against the oracle's own run-to-run agreement (16-bit rounding makes a random-init
    ResNet at batch 8 chaotic: see test_gpu_engine.py::test_default_mode_differs_from_deterministic_only_by_rounding_noise)."""
    import copy
    from pytorch_distributed_b200.models import resnet50
    from pytorch_distributed_b200.nn import CrossEntropyLoss
    from pytorch_distributed_b200.parallel import DataParallel
    monkeypatch.setenv("PDB200_DP_SHARD_MIN", shard_min)       # "2": the sharded replicate / reduce on two GPUs
    torch.manual_seed(0)
    base = resnet50().cuda(0)
    ref = copy.deepcopy(base)
    base.enable_native(batch_size=8, dtype=dt, deterministic=det)
    ref.enable_native(batch_size=8, dtype=dt, deterministic=det)
    dp = DataParallel(base, device_ids=[0, 1], output_device=0)
    x = torch.randn(16, 3, 128, 128, device="cuda:0")
    y = torch.randint(0, 1000, (16,), device="cuda:0")
    # the batch comes from pinned host memory, each half straight to its GPU (what harness.train does in dp mode)
    out = dp(dp.prescatter(x.cpu().pin_memory()))
    assert out.shape == (16, 1000) and out.device.index == 0
    (CrossEntropyLoss()(out, y) * 2 * 64).backward()          # mean over 16 rows (x64: keep fp16 grads normal)
    torch.cuda.synchronize()

    def sequential():
        ref._pdb200_arena.grads.zero_()
        for cx, cy in zip(x.chunk(2), y.chunk(2)):
            (CrossEntropyLoss()(ref(cx), cy) * 64).backward()   # mean over 8 rows each -> same total scale
        torch.cuda.synchronize()
        return ref._pdb200_arena.grads.clone()

    cos = torch.nn.functional.cosine_similarity
    r1, r2 = sequential(), sequential()
    noise = float(cos(r1, r2, dim=0))
    a = base._pdb200_arena.grads
    got = float(cos(a, r1, dim=0))
    print(f"[{dt}, deterministic={det}] DP vs sequential cos {got:.6f}; sequential run-to-run cos {noise:.6f}")
    if det:
        assert torch.equal(r1, r2)                              # the oracle is bit-reproducible
        assert got > 0.99999, got
        assert float((a - r1).abs().max() / r1.abs().max()) < 1e-5
    else:
        assert got > noise - 0.05, (got, noise)
        assert 0.9 < float(a.norm() / r1.norm()) < 1.1
    # replica parameters follow the root on the next forward (K4b)
    with torch.no_grad():
        base.fc.bias.add_(1.0)
    dp(x)
    torch.cuda.synchronize()
    assert torch.equal(dp._replicas[1].fc.bias.cpu(), base.fc.bias.cpu())
