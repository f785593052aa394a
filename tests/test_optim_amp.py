import copy

import torch

from pytorch_distributed_b200 import amp
from pytorch_distributed_b200.models import ResNet
from pytorch_distributed_b200.optim import SGD, StepLR
from pytorch_distributed_b200.utils import ParamArena


def _tiny():
    torch.manual_seed(0)
    return ResNet((1, 1, 1, 1), num_classes=10, width=8)


def test_sgd_arena_matches_torch():
    a, b = _tiny(), None
    b = copy.deepcopy(a)
    ParamArena(a)
    # small lr: the tiny random net is chaotic at lr 0.1 and would amplify 1e-6 conv-algorithm noise
    oa = SGD(a.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    ob = torch.optim.SGD(b.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    assert oa.arena is not None
    x = torch.randn(4, 3, 32, 32)
    for _ in range(2):
        for m, o in ((a, oa), (b, ob)):
            o.zero_grad()
            m(x).square().mean().backward()
            o.step()
    for (n, p), q in zip(a.named_parameters(), b.parameters()):
        assert torch.allclose(p, q, rtol=1e-3, atol=1e-4), n
    # grads still alias the arena after zero_grad(set_to_none=True)
    assert a.conv1.weight.grad.data_ptr() == oa.arena.grads.data_ptr()


def test_sgd_state_dict_interchange():
    a = _tiny()
    b = copy.deepcopy(a)
    ParamArena(a)
    oa = SGD(a.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    ob = torch.optim.SGD(b.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    x = torch.randn(2, 3, 32, 32)
    for m, o in ((a, oa), (b, ob)):
        m(x).sum().backward()
        o.step()
    sa, sb = oa.state_dict(), ob.state_dict()
    assert sa["param_groups"][0].keys() == sb["param_groups"][0].keys()
    assert sa["param_groups"][0]["params"] == sb["param_groups"][0]["params"]
    assert set(sa["state"].keys()) == set(sb["state"].keys())
    ob.load_state_dict(sa)                      # ours -> torch
    oa.load_state_dict(sb)                      # torch -> ours, lands in the momentum arena
    p0 = oa.arena.parameters()[0]
    assert oa.state[p0]["momentum_buffer"].untyped_storage().data_ptr() == oa.arena.momentum.untyped_storage().data_ptr()


def test_steplr_matches_torch():
    m = torch.nn.Linear(2, 2)
    oa = torch.optim.SGD(m.parameters(), lr=0.1)
    ob = torch.optim.SGD(m.parameters(), lr=0.1)
    a, b = StepLR(oa, 30, 0.1), torch.optim.lr_scheduler.StepLR(ob, 30, 0.1)
    assert a.state_dict().keys() == b.state_dict().keys()
    for _ in range(95):
        oa.step(), ob.step(), a.step(), b.step()
        assert abs(oa.param_groups[0]["lr"] - ob.param_groups[0]["lr"]) < 1e-12
    b.load_state_dict(a.state_dict())
    a.load_state_dict(b.state_dict())
    assert a.last_epoch == 95


def test_grad_scaler_state_machine():
    m = _tiny()
    ParamArena(m)
    opt = SGD(m.parameters(), lr=0.1, momentum=0.9)
    s = amp.GradScaler(init_scale=1024.0, growth_interval=2)
    x = torch.randn(2, 3, 32, 32)
    before = opt.arena.params.clone()
    loss = m(x).mean()
    s.scale(loss).backward()
    opt.arena.grads[5] = float("inf")             # overflow: step must be skipped, scale halves
    s.step(opt), s.update(), opt.zero_grad()
    assert torch.equal(before, opt.arena.params) and s.get_scale() == 512.0
    for i in range(2):                            # two clean steps -> growth
        s.scale(m(x).mean()).backward()
        s.step(opt), s.update(), opt.zero_grad()
    assert s.get_scale() == 1024.0 and not torch.equal(before, opt.arena.params)
    sd = s.state_dict()
    assert set(sd) == set(torch.amp.GradScaler("cpu").state_dict())
    t = amp.GradScaler()
    t.load_state_dict(sd)
    assert t.get_scale() == 1024.0


def test_grad_scaler_unscale_value():
    m = _tiny()
    ParamArena(m)
    ref = copy.deepcopy(m)
    opt = SGD(m.parameters(), lr=0.1)
    opt_ref = torch.optim.SGD(ref.parameters(), lr=0.1)
    s = amp.GradScaler(init_scale=256.0)
    x = torch.randn(2, 3, 32, 32)
    s.scale(m(x).mean()).backward()
    s.step(opt), s.update()
    ref(x).mean().backward()
    opt_ref.step()
    for p, q in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(p, q, rtol=1e-4, atol=1e-6)


def test_sgd_on_a_subset_of_an_arena_touches_nothing_else():
    """SGD(model.fc.parameters()) on an arena-backed model must behave like torch.optim.SGD: the fused whole-arena
    step is only taken when the optimizer owns every parameter of the arena."""
    import torch
    from pytorch_distributed_b200.optim import SGD
    from pytorch_distributed_b200.utils.arena import ParamArena
    torch.manual_seed(0)
    m = torch.nn.Sequential(torch.nn.Linear(8, 8), torch.nn.Linear(8, 4))
    ParamArena(m, torch.device("cpu"))
    opt = SGD(m[1].parameters(), lr=0.1, momentum=0.9, weight_decay=0.1)
    assert opt.arena is None
    before0 = m[0].weight.detach().clone()
    before1 = m[1].weight.detach().clone()
    m(torch.randn(2, 8)).sum().backward()
    opt.step()
    assert torch.equal(m[0].weight, before0) and not torch.equal(m[1].weight, before1)
    full = SGD(m.parameters(), lr=0.1)
    assert full.arena is m._pdb200_arena
    m[0].bias.requires_grad_(False)
    assert SGD(m.parameters(), lr=0.1).arena is None          # a frozen parameter: no whole-arena step either


def test_grad_scaler_resume_keeps_the_growth_counter():
    """save -> new scaler -> load (before the first scale(), as harness.main does) -> the tracker survives."""
    import torch
    from pytorch_distributed_b200 import amp
    a = amp.GradScaler(init_scale=256.0, growth_interval=5)
    a.scale(torch.ones(()))
    for _ in range(3):
        a.update()
    sd = a.state_dict()
    assert sd["_growth_tracker"] == 3
    b = amp.GradScaler()
    b.load_state_dict(sd)
    b.scale(torch.ones(()))                                    # lazy init happens here
    assert b.state_dict()["_growth_tracker"] == 3 and b.get_scale() == 256.0
    b.update(), b.update()
    assert b.get_scale() == 512.0                              # grows on schedule, as the uninterrupted run would
