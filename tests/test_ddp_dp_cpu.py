import copy

import pytest
import torch

from pytorch_distributed_b200 import launcher
from pytorch_distributed_b200.models import ResNet
from pytorch_distributed_b200.parallel import DataParallel
from pytorch_distributed_b200.parallel.scatter_gather import gather, scatter

from . import _mp_helpers as H


@pytest.mark.timeout(600)
def test_ddp_gloo_matches_torch_ddp(tmp_path, free_port):
    world, steps = 2, 3
    launcher.spawn(H.ddp_worker, args=(world, free_port, str(tmp_path), True, steps), nprocs=world)
    launcher.spawn(H.ddp_worker, args=(world, free_port + 1, str(tmp_path), False, steps), nprocs=world)
    ours = [torch.load(tmp_path / f"ours_{r}.pt") for r in range(world)]
    ref = [torch.load(tmp_path / f"torch_{r}.pt") for r in range(world)]
    assert torch.equal(ours[0]["params"], ours[1]["params"])             # replicas stay bit-identical
    for r in range(world):
        assert ours[r]["losses"] == pytest.approx(ref[r]["losses"], rel=2e-3, abs=1e-4)
    assert torch.allclose(ours[0]["params"], ref[0]["params"], rtol=5e-3, atol=5e-4)
    # a user-supplied Python comm hook (hook(state, bucket) -> Future) replaces the built-in reduction
    launcher.spawn(H.ddp_worker, args=(world, free_port + 2, str(tmp_path), True, steps, False, "py"), nprocs=world)
    hooked = [torch.load(tmp_path / f"ours_py_{r}.pt") for r in range(world)]
    assert torch.equal(hooked[0]["params"], hooked[1]["params"])
    assert torch.allclose(hooked[0]["params"], ours[0]["params"], rtol=1e-4, atol=1e-6)


@pytest.mark.timeout(600)
def test_syncbn_gloo_matches_torch(tmp_path, free_port):
    world, steps = 2, 2
    launcher.spawn(H.ddp_worker, args=(world, free_port, str(tmp_path), True, steps, True), nprocs=world)
    ours = [torch.load(tmp_path / f"ours_{r}.pt") for r in range(world)]
    assert torch.equal(ours[0]["params"], ours[1]["params"])
    # single-process oracle: plain BN over the concatenated batch == SyncBN over the shards
    torch.manual_seed(1234)
    m = ResNet((1, 1, 1, 1), num_classes=10, width=8)
    opt = torch.optim.SGD(m.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator().manual_seed(99)
    xs = torch.randn(steps, world, 4, 3, 32, 32, generator=g)
    ys = torch.randint(0, 10, (steps, world, 4), generator=g)
    for s in range(steps):
        opt.zero_grad()
        torch.nn.functional.cross_entropy(m(xs[s].flatten(0, 1)), ys[s].flatten()).backward()
        opt.step()
    flat = torch.cat([p.detach().reshape(-1) for p in m.parameters()])
    assert torch.allclose(ours[0]["params"], flat, rtol=5e-3, atol=5e-4)


def test_scatter_gather_autograd():
    x = torch.randn(7, 3, requires_grad=True)
    devs = [torch.device("cpu")] * 3
    parts = scatter((x,), devs)
    assert [p[0].shape[0] for p in parts] == [3, 3, 1]
    out = gather([p[0] * (i + 1) for i, p in enumerate(parts)], torch.device("cpu"))
    out.sum().backward()
    assert torch.equal(x.grad, torch.tensor([1.] * 3 + [2.] * 3 + [3.]).view(7, 1).expand(7, 3))
    d = scatter(({"a": x, "b": 5},), devs)
    assert d[1][0]["b"] == 5 and d[1][0]["a"].shape[0] == 3


def test_data_parallel_sums_replica_grads():
    torch.manual_seed(0)
    base = ResNet((1, 1, 1, 1), num_classes=10, width=8)
    ref = copy.deepcopy(base)
    dp = DataParallel(base, device_ids=["cpu", "cpu"])
    x, y = torch.randn(8, 3, 32, 32), torch.randint(0, 10, (8,))
    out = dp(x)
    assert out.shape == (8, 10)
    torch.nn.functional.cross_entropy(out, y, reduction="sum").backward()
    # oracle: per-replica BN statistics => run each chunk separately and SUM the gradients
    for chunk_x, chunk_y in zip(x.chunk(2), y.chunk(2)):
        torch.nn.functional.cross_entropy(ref(chunk_x), chunk_y, reduction="sum").backward()
    for (n, p), q in zip(base.named_parameters(), ref.parameters()):
        assert torch.allclose(p.grad, q.grad, rtol=1e-3, atol=1e-3 * float(q.grad.abs().max())), n
    # replica 0 is the module itself: its running stats persist (first chunk's statistics)
    r0 = copy.deepcopy(ref)
    assert dp._replicas[0] is base
    # second forward re-broadcasts parameters: replicas follow an update of the root
    with torch.no_grad():
        base.fc.bias.add_(1.0)
    dp(x)
    assert torch.equal(dp._replicas[1].fc.bias, base.fc.bias)


def test_data_parallel_module_on_wrong_device_message():
    m = ResNet((1, 1, 1, 1), num_classes=10, width=8)
    dp = DataParallel(m, device_ids=["cpu"])
    assert dp(torch.randn(2, 3, 32, 32)).shape == (2, 10)


def test_data_parallel_balance_warning(monkeypatch):
    """min/max memory or SM count below 0.75 across the replicas' GPUs warns (torch:nn/parallel/data_parallel.py:24-48)."""
    import types
    import warnings
    from pytorch_distributed_b200.parallel import data_parallel as dpm
    props = {0: types.SimpleNamespace(total_memory=180 << 30, multi_processor_count=148),
             1: types.SimpleNamespace(total_memory=80 << 30, multi_processor_count=132)}
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(torch.cuda, "get_device_properties", lambda d: props[torch.device(d).index])
    devs = [torch.device("cuda", 0), torch.device("cuda", 1)]
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        msg = dpm.check_balance(devs)
    assert msg and "memory" in msg and len(w) == 1
    props[1] = props[0]
    assert dpm.check_balance(devs) is None
    assert dpm.check_balance([torch.device("cpu"), torch.device("cpu")]) is None


def test_data_parallel_accepts_a_prescattered_batch():
    """`prescatter` splits like `scatter` does; forward on the pre-split batch gives the same output as on the tensor."""
    torch.manual_seed(0)
    m = ResNet((1, 1, 1, 1), num_classes=10, width=8)
    dp = DataParallel(m, device_ids=["cpu", "cpu", "cpu"])
    x = torch.randn(7, 3, 32, 32)                       # 3 + 3 + 1: uneven chunks
    pre = dp.prescatter(x)
    assert [c.size(0) for c in pre.chunks] == [3, 3, 1] and pre.size(0) == 7 and tuple(pre.shape) == tuple(x.shape)
    dp.eval()
    with torch.no_grad():
        assert torch.allclose(dp(pre), dp(x), rtol=1e-5, atol=1e-6)
