"""Native step graph vs the stock fp32 model: logits, loss, every parameter gradient, BN buffers."""
import copy

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from pytorch_distributed_b200 import amp, harness  # noqa: E402
from pytorch_distributed_b200.config import TrainConfig  # noqa: E402
from pytorch_distributed_b200.models import resnet50  # noqa: E402
from pytorch_distributed_b200.nn import CrossEntropyLoss  # noqa: E402
from pytorch_distributed_b200.optim import SGD  # noqa: E402


def _cos(a, b):
    a, b = a.float().flatten(), b.float().flatten()
    return float(F.cosine_similarity(a, b, dim=0))


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
def test_forward_backward_matches_fp32_model(dt):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)
    dev = torch.device("cuda")
    ref = resnet50().to(dev)
    ours = copy.deepcopy(ref)
    ours.enable_native(batch_size=16, dtype=dt)
    x = torch.randn(16, 3, 96, 96, device=dev)
    y = torch.randint(0, 1000, (16,), device=dev)

    ref.train()
    out_ref = ref(x)
    loss_ref = F.cross_entropy(out_ref, y)
    loss_ref.backward()

    ours.train()
    out = ours(x)
    assert out.dtype == dt and out.shape == (16, 1000)
    loss = CrossEntropyLoss()(out, y)
    loss.backward()
    torch.cuda.synchronize()

    # yardstick: the same stock model under torch.autocast in the same 16-bit format — 16-bit rounding
    # through 53 conv+BN layers is what separates us from fp32, so we must be about as close as it is
    yard = copy.deepcopy(ref)
    for p in yard.parameters():
        p.grad = None
    with torch.autocast("cuda", dtype=dt):
        out_y = yard(x)
        loss_y = F.cross_entropy(out_y.float(), y)
    loss_y.backward()
    cos_out, cos_yard = _cos(out, out_ref), _cos(out_y, out_ref)
    print(f"[{dt}] logits cos: ours {cos_out:.5f}  autocast {cos_yard:.5f}; loss ours {float(loss):.4f} "
          f"autocast {float(loss_y):.4f} fp32 {float(loss_ref):.4f}")
    assert cos_out > min(0.995, cos_yard - 0.02), (cos_out, cos_yard)
    assert abs(float(loss) - float(loss_ref)) < 0.05 * abs(float(loss_ref))
    worst, worst_yard = (1.0, None), (1.0, None)
    ga, gr, gy = [], [], []
    for (n, p), q, r in zip(ours.named_parameters(), ref.parameters(), yard.parameters()):
        assert p.grad.data_ptr() == ours._pdb200_arena.grad_view(n).data_ptr()
        c, cy = _cos(p.grad, q.grad), _cos(r.grad, q.grad)
        if c < worst[0]:
            worst = (c, n)
        if cy < worst_yard[0]:
            worst_yard = (cy, n)
        # Some gradients (BN gammas of early layers at this tiny batch) are rounding noise in ANY 16-bit run:
        # only parameters that the stock autocast run itself gets right are compared one by one
        if cy > 0.9:
            assert c > cy - 0.05, (n, c, cy)
            rel = float(p.grad.float().norm() / (q.grad.norm() + 1e-20))
            assert 0.7 < rel < 1.4, (n, rel)
        ga.append(p.grad.float().flatten()), gr.append(q.grad.flatten()), gy.append(r.grad.float().flatten())
    ga, gr, gy = torch.cat(ga), torch.cat(gr), torch.cat(gy)
    print(f"[{dt}] whole-gradient cos: ours {_cos(ga, gr):.4f}  autocast {_cos(gy, gr):.4f}; "
          f"worst per-param: ours {worst}  autocast {worst_yard}")
    assert _cos(ga, gr) > _cos(gy, gr) - 0.03
    # BN running statistics were updated like torch's
    for (n, b), c in zip(ours.named_buffers(), ref.buffers()):
        if b.is_floating_point():
            # deep-layer statistics carry the accumulated 16-bit rounding of everything before them
            assert _cos(b, c) > 0.98 and 0.8 < float(b.norm() / c.norm()) < 1.25, n
        else:
            assert int(b) == int(c) == 1, n
    # eval forward uses running stats
    ours.eval(), ref.eval()
    with torch.no_grad():
        assert _cos(ours(x), ref(x)) > 0.99


def test_training_steps_match_torch_engine():
    """Optimisation steps through the public harness step: native vs the stock-op engine."""
    torch.manual_seed(1)
    dev = torch.device("cuda")
    # fp16 + loss scaling is the reference's "apex" configuration and is far less noisy than bf16 at this
    # tiny batch (logits cos 0.999 vs 0.97), so trajectories can be compared meaningfully
    cfg = TrainConfig(mode="ddp_amp", batch_size=32, dtype="fp16", device="cuda")
    a = harness.build_model(cfg, dev)
    b = copy.deepcopy(a)
    a.enable_native(batch_size=32, dtype=torch.float16)
    arena_a, arena_b = harness.ensure_arenas(a, dev), harness.ensure_arenas(b, dev)
    x = torch.randn(32, 3, 64, 64, device=dev)
    y = torch.randint(0, 1000, (32,), device=dev)
    losses, deltas = [], []
    for m, arena, stock in ((a, arena_a, False), (b, arena_b, True)):
        opt = SGD(m.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
        sc = amp.GradScaler(init_scale=1024.0, device=dev)     # no fp16 overflow skips in either engine
        crit = CrossEntropyLoss()
        m.train()
        before = arena.params.clone()
        ls = [float(harness.train_step(m, crit, opt, x, y, cfg, dev, sc, stock_autocast=stock))]
        deltas.append(arena.params - before)
        ls += [float(harness.train_step(m, crit, opt, x, y, cfg, dev, sc, stock_autocast=stock)) for _ in range(9)]
        losses.append(ls)
    print("native", [round(v, 3) for v in losses[0]], "stock", [round(v, 3) for v in losses[1]])
    # the first update (whole arena, dominated by the conv weights) points the same way
    print("first-update cos", _cos(deltas[0], deltas[1]))
    assert _cos(deltas[0], deltas[1]) > 0.8, _cos(deltas[0], deltas[1])
    assert 0.8 < float(deltas[0].norm() / deltas[1].norm()) < 1.25
    assert abs(losses[0][0] - losses[1][0]) < 0.05 * losses[1][0]
    assert abs(losses[0][1] - losses[1][1]) < 0.08 * losses[1][1]
    # both engines fit the fixed batch at a similar pace
    assert losses[0][-1] < losses[0][0] and abs(losses[0][-1] - losses[1][-1]) < 0.25 * losses[1][0], losses


def test_tf32_engine_matches_fp32_model():
    """fp32 storage + kind::tf32 GEMMs (what `single` / `dp` / `ddp` run by default) against the stock model in true
    fp32; yardstick = the stock model with cuDNN/cuBLAS TF32 enabled, i.e. what the reference's fp32 scripts execute."""
    torch.manual_seed(0)
    dev = torch.device("cuda")
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    ref = resnet50().to(dev)
    ours = copy.deepcopy(ref)
    ours.enable_native(batch_size=16, dtype=torch.float32)
    assert ours._pdb200_arena.shadow is None          # the master weights ARE the GEMM operands
    x = torch.randn(16, 3, 96, 96, device=dev)
    y = torch.randint(0, 1000, (16,), device=dev)
    ref.train()
    out_ref = ref(x)
    loss_ref = F.cross_entropy(out_ref, y)
    loss_ref.backward()
    ours.train()
    out = ours(x)
    assert out.dtype == torch.float32
    loss = CrossEntropyLoss()(out, y)
    loss.backward()
    yard = copy.deepcopy(ref)
    for p in yard.parameters():
        p.grad = None
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        out_y = yard(x)
        F.cross_entropy(out_y, y).backward()
    finally:
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
    torch.cuda.synchronize()
    cos_out, cos_yard = _cos(out, out_ref), _cos(out_y, out_ref)
    print(f"[tf32] logits cos: ours {cos_out:.6f}  stock-tf32 {cos_yard:.6f}; loss ours {float(loss):.5f} fp32 {float(loss_ref):.5f}")
    assert cos_out > min(0.9995, cos_yard - 1e-3), (cos_out, cos_yard)
    assert abs(float(loss) - float(loss_ref)) < 5e-3 * abs(float(loss_ref))
    ga, gr, gy = [], [], []
    for (n, p), q, r in zip(ours.named_parameters(), ref.parameters(), yard.parameters()):
        c, cy = _cos(p.grad, q.grad), _cos(r.grad, q.grad)
        if cy > 0.9:
            assert c > cy - 0.03, (n, c, cy)
            rel = float(p.grad.norm() / (q.grad.norm() + 1e-20))
            assert 0.9 < rel < 1.1, (n, rel)
        ga.append(p.grad.flatten()), gr.append(q.grad.flatten()), gy.append(r.grad.flatten())
    ga, gr, gy = torch.cat(ga), torch.cat(gr), torch.cat(gy)
    print(f"[tf32] whole-gradient cos: ours {_cos(ga, gr):.5f}  stock-tf32 {_cos(gy, gr):.5f}")
    assert _cos(ga, gr) > min(0.995, _cos(gy, gr) - 0.01)
    for (n, b), c in zip(ours.named_buffers(), ref.buffers()):
        if b.is_floating_point():
            assert _cos(b, c) > 0.9999 and 0.99 < float(b.norm() / c.norm()) < 1.01, n
    ours.eval(), ref.eval()
    with torch.no_grad():
        assert _cos(ours(x), ref(x)) > 0.9995


@pytest.mark.parametrize("engine", ["native", "torch"])
def test_fp32_single_mode_trains_through_the_harness(engine, tmp_path):
    """`resnet_single_gpu.py` with no flags = fp32: on the native engine (TF32 GEMMs) and on the stock-op engine
    (fp32 logits into the fused loss — used to raise a TypeError)."""
    cfg = TrainConfig(mode="single", batch_size=8, epochs=1, max_steps=3, max_val_steps=1, image_size=64,
                      train_samples=64, val_samples=16, engine=engine, device="cuda", save_path=str(tmp_path))
    assert cfg.resolved_dtype() == "fp32"
    acc = harness.main(cfg)
    assert 0.0 <= acc <= 1.0


def _grads_of(model, x, y, scale=1.0):
    model._pdb200_arena.grads.zero_()
    out = model(x)
    (CrossEntropyLoss()(out, y) * scale).backward()
    torch.cuda.synchronize()
    return out.detach().clone(), model._pdb200_arena.grads.clone()


@pytest.mark.parametrize("dt", [torch.float16, torch.float32])
def test_deterministic_mode_is_bitwise_reproducible(dt):
    """PDB200_DETERMINISTIC: same input, same bits — across eager runs AND the CUDA-graph replays that follow.  A race
    anywhere in the step (pipeline barriers, TMA stores vs later reads, the statistics fold) would show up as a
    difference here; this is the racecheck the GPU pool's closed compute-sanitizer cannot provide."""
    torch.manual_seed(3)
    dev = torch.device("cuda")
    m = resnet50().to(dev)
    m.enable_native(batch_size=8, dtype=dt, deterministic=True)
    m.train()
    x = torch.randn(8, 3, 128, 128, device=dev)
    y = torch.randint(0, 1000, (8,), device=dev)
    rm0 = {n: b.clone() for n, b in m.named_buffers()}
    runs = []
    for _ in range(5):                       # steps 1-2 eager, 3 captures, 4-5 replay the graphs
        for n, b in m.named_buffers():
            b.copy_(rm0[n])
        runs.append(_grads_of(m, x, y, 64.0))
    assert m._engine._plans and all(p.fwd_graph is not None and p.bwd_graphs is not None for p in m._engine._plans.values())
    for out, g in runs[1:]:
        assert torch.equal(out, runs[0][0]) and torch.equal(g, runs[0][1])


def test_default_mode_differs_from_deterministic_only_by_rounding_noise():
    """Where the 0.89 run-to-run gradient cosine of round 1's fp16 batch-8 oracle came from.  The deterministic engine
    reproduces itself bit for bit (test above), so nothing races; the default engine differs from it only in the ORDER
    of fp32 additions (atomics, split-K), i.e. in the last bit of a few statistics.  But every activation is rounded
    to a 10-bit mantissa on its way into the tensor cores (fp16 storage, or the TF32 conversion of fp32 storage), so a
    last-bit difference flips roundings downstream and after a few layers the two runs carry two independent
    realisations of the 2^-11 quantisation noise.  How far apart that puts two gradients is a property of the test
    case, measured here on STOCK PyTorch: fp32 vs TF32 / fp16-autocast of the same model and batch.  A random-init
    ResNet-50 at batch 8 x 128^2 (layer4 BatchNorm sees 128 samples) is that sensitive; the benched shape is not
    (test_benched_shape_step_matches_stock_engine)."""
    torch.manual_seed(4)
    dev = torch.device("cuda")
    x = torch.randn(8, 3, 128, 128, device=dev)
    y = torch.randint(0, 1000, (8,), device=dev)
    base = resnet50().to(dev)

    def stock(mode):
        m = copy.deepcopy(base).train()
        torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = mode == "tf32"
        try:
            if mode == "fp16":
                with torch.autocast("cuda", dtype=torch.float16):
                    loss = F.cross_entropy(m(x).float(), y)
            else:
                loss = F.cross_entropy(m(x), y)
            (loss * 64.0).backward()
        finally:
            torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
        return torch.cat([p.grad.float().flatten() for p in m.parameters()])

    g32 = stock("fp32")
    yard = {torch.float32: _cos(stock("tf32"), g32), torch.float16: _cos(stock("fp16"), g32)}
    for dt in (torch.float32, torch.float16):
        a, b = copy.deepcopy(base), copy.deepcopy(base)
        a.enable_native(batch_size=8, dtype=dt, deterministic=True)
        b.enable_native(batch_size=8, dtype=dt, deterministic=False)
        a.train(), b.train()
        (_, ga), (_, gb), (_, gb2) = _grads_of(a, x, y, 64.0), _grads_of(b, x, y, 64.0), _grads_of(b, x, y, 64.0)
        c_dd, c_rr = _cos(ga, gb), _cos(gb, gb2)
        # arena order == parameter order, but the arena pads between tensors: compare through the views
        gd = torch.cat([p.grad.float().flatten() for p in a.parameters()])
        c_truth = _cos(gd, g32)
        print(f"[{dt}] default vs deterministic cos {c_dd:.4f}; default run-to-run {c_rr:.4f}; deterministic vs stock fp32 "
              f"{c_truth:.4f}; stock reduced-precision vs stock fp32 (sensitivity of this case) {yard[dt]:.4f}")
        assert c_dd > yard[dt] - 0.08 and c_rr > yard[dt] - 0.08, (dt, c_dd, c_rr, yard[dt])
        assert c_truth > yard[dt] - 0.08, (dt, c_truth, yard[dt])


def _group_of(name: str) -> str:
    if name.startswith("fc."):
        return name
    stage = name.split(".")[0] if name.startswith("layer") else "stem"
    kind = "conv" if ("conv" in name or "downsample.0" in name) else ("gamma" if name.endswith("weight") else "beta")
    return f"{stage}.{kind}"


def _stock_step(base, x, y, mode, scale=1.0):
    """Loss, logits and per-parameter gradients of the stock model in `mode` (fp32 | tf32 | torch dtype for autocast);
    everything else it allocated is released before returning (batch 400 activations are tens of GB)."""
    import gc
    m = copy.deepcopy(base).train()
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = mode == "tf32"
    try:
        if isinstance(mode, torch.dtype):
            with torch.autocast("cuda", dtype=mode):
                out = m(x)
                loss = F.cross_entropy(out.float(), y)
        else:
            out = m(x)
            loss = F.cross_entropy(out, y)
        (loss * scale).backward()
    finally:
        torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    res = float(loss), out.detach().float().clone(), [p.grad.float().flatten() / scale for p in m.parameters()]
    del m, out, loss
    gc.collect()
    torch.cuda.empty_cache()
    return res


@pytest.mark.timeout(900)
@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16, torch.float32])
def test_benched_shape_step_matches_stock_engine(dt):
    """One whole step at the shape bench.py times (batch 400, 224x224) — where the CTA-pair tiles, the one-wave split-K
    heuristic and the replica folds are the variants actually selected — against the stock fp32 model: logits, loss and
    the gradient norm / direction of every parameter group (stage x {conv, gamma, beta}, fc).  The gradient of a
    random-init ResNet-50 on random labels is mostly rounding noise below layer4 in ANY reduced precision (stock bf16
    autocast: cos 0.12 with fp32 for the stem), so directions are held to the stock model run in the same precision on
    the same batch, and only where that run itself is meaningful; norms are compared everywhere.  The sharp per-kernel
    check at these shapes is tests/test_gpu_conv.py::test_benched_shapes."""
    import gc
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    gc.collect()
    torch.cuda.empty_cache()
    torch.manual_seed(5)
    dev = torch.device("cuda")
    base = resnet50().to(dev)
    x = torch.randn(400, 3, 224, 224, device=dev)
    y = torch.randint(0, 1000, (400,), device=dev)
    loss_ref, out_ref, g_ref = _stock_step(base, x, y, "fp32")
    _, _, g_yard = _stock_step(base, x, y, "tf32" if dt == torch.float32 else dt, 1.0 if dt == torch.float32 else 1024.0)
    ours = copy.deepcopy(base)
    ours.enable_native(batch_size=400, dtype=dt)
    ours.train()
    out = ours(x)
    loss = CrossEntropyLoss()(out, y)
    (loss * 1024.0).backward()                      # a loss scale, as in the amp step (keeps fp16 gradients normal)
    torch.cuda.synchronize()
    cos_out = _cos(out, out_ref)
    tol_cos, tol_loss, tol_norm = {torch.float32: (0.9995, 2e-3, 0.03), torch.float16: (0.995, 1e-2, 0.1),
                                   torch.bfloat16: (0.97, 3e-2, 0.15)}[dt]
    print(f"[{dt} @400x224] logits cos {cos_out:.5f}; loss {float(loss):.4f} vs fp32 {loss_ref:.4f}")
    assert cos_out > tol_cos, cos_out
    assert abs(float(loss) - loss_ref) < tol_loss * loss_ref
    groups = {}
    for (n, p), q, r in zip(ours.named_parameters(), g_ref, g_yard):
        g = groups.setdefault(_group_of(n), [[], [], []])
        g[0].append(p.grad.float().flatten() / 1024.0), g[1].append(q), g[2].append(r)
    bad = []
    for name, (a, b, c3) in groups.items():
        a, b, c3 = torch.cat(a), torch.cat(b), torch.cat(c3)
        ratio, c, cy, ry = float(a.norm() / (b.norm() + 1e-20)), _cos(a, b), _cos(c3, b), float(c3.norm() / (b.norm() + 1e-20))
        print(f"[{dt} @400x224] {name:14s} |g|/|g_fp32| {ratio:.4f} (stock {ry:.4f})  cos {c:.5f} (stock {cy:.5f})  "
              f"cos(ours, stock same precision) {_cos(a, c3):.5f}")
        # a group whose stock reduced-precision gradient is itself partly noise (cos < 0.9 with fp32) has a noisy norm too
        ntol = max(tol_norm, 1.5 * abs(ry - 1.0)) if cy > 0.9 else 0.4
        if abs(ratio - 1.0) > ntol or (cy > 0.9 and c < cy - 0.03):
            bad.append((name, ratio, c, ry, cy))
    assert not bad, bad
    del ours, out, loss, groups
    gc.collect()
    torch.cuda.empty_cache()
