import torch
import torchvision

from pytorch_distributed_b200 import checkpoint
from pytorch_distributed_b200.models import ResNet, resnet50
from pytorch_distributed_b200.optim import SGD, StepLR
from pytorch_distributed_b200.utils import BufferArena, ParamArena


def test_state_dict_is_torchvision_resnet50():
    ours, tv = resnet50().state_dict(), torchvision.models.resnet50().state_dict()
    assert len(ours) == 320
    assert [(k, v.shape, v.dtype) for k, v in ours.items()] == [(k, v.shape, v.dtype) for k, v in tv.items()]


def test_param_count_and_buckets():
    m = resnet50()
    assert sum(p.numel() for p in m.parameters()) == 25_557_032 and len(list(m.parameters())) == 161
    a = ParamArena(m)
    mb = [round((e - s) * 4 / 1e6, 2) for s, e in a.bucket_ranges()]
    assert mb == [8.2, 31.5, 26.26, 26.55, 9.72]             # SURVEY §2.4 K1 bucket sizes


def test_forward_parity_with_torchvision_and_arena():
    torch.manual_seed(0)
    m, tv = resnet50().eval(), torchvision.models.resnet50().eval()
    tv.load_state_dict(m.state_dict())
    x = torch.randn(2, 3, 64, 64)
    ref = tv(x)
    assert torch.allclose(m(x), ref, atol=1e-4)
    ParamArena(m), BufferArena(m)                             # re-homing must not change the function
    assert torch.allclose(m(x), ref, atol=1e-3)
    m.load_state_dict(tv.state_dict())                        # torchvision -> ours through the views
    assert m.conv1.weight.data_ptr() == m._pdb200_arena.params.data_ptr()
    assert m.bn1.running_mean.data_ptr() == m._pdb200_buffer_arena.floats.data_ptr()


def test_latest_and_best_roundtrip(tmp_path):
    torch.manual_seed(0)
    m = ResNet((1, 1, 1, 1), num_classes=10, width=8)
    ParamArena(m), BufferArena(m)
    opt = SGD(m.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    sch = StepLR(opt, 30, 0.1)
    m(torch.randn(2, 3, 32, 32)).mean().backward()
    opt.step()
    checkpoint.save_latest(tmp_path, m, opt, sch, 0.25, 3, 17)
    raw = torch.load(tmp_path / "latest.pt", map_location="cpu", weights_only=False)
    assert set(raw) == {"model", "optimizer", "scheduler", "acc", "epoch", "step"}
    assert not any(k.startswith("module.") for k in raw["model"])
    assert all(v.is_contiguous() for v in raw["model"].values())
    m2 = ResNet((1, 1, 1, 1), num_classes=10, width=8)
    ParamArena(m2), BufferArena(m2)
    opt2 = SGD(m2.parameters(), lr=0.5, momentum=0.9, weight_decay=1e-4)
    sch2 = StepLR(opt2, 30, 0.1)
    assert checkpoint.load_latest(tmp_path, m2, opt2, sch2) == (0.25, 3, 17)
    for p, q in zip(m.parameters(), m2.parameters()):
        assert torch.equal(p, q)
    assert torch.equal(opt.arena.momentum, opt2.arena.momentum) and opt2.param_groups[0]["lr"] == 0.1
    # the same file loads into stock torch objects (format parity)
    so = torch.optim.SGD(ResNet((1, 1, 1, 1), num_classes=10, width=8).parameters(), lr=0.1, momentum=0.9)
    so.load_state_dict(raw["optimizer"])
    checkpoint.save_best(tmp_path, m)
    best = torch.load(tmp_path / "best.pt", map_location="cpu")
    assert list(best) == list(m.state_dict())
