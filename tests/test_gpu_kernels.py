"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference of the same op."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from pytorch_distributed_b200 import ops  # noqa: E402
from pytorch_distributed_b200.optim.sgd import _flat_sgd  # noqa: E402

from ._gpu_helpers import rel_err  # noqa: E402

DT = [torch.bfloat16, torch.float16, torch.float32]      # fp32 storage = the TF32 path of the three fp32 modes


def test_library_loaded():
    assert ops.available(), "libpdb200.so must load on a GPU box (no silent fallback)"


@pytest.mark.parametrize("first", [True, False])
@pytest.mark.parametrize("nesterov", [False, True])
def test_sgd_step(first, nesterov):
    n = 1 << 20
    g = torch.Generator(device="cuda").manual_seed(0)
    p = torch.randn(n, device="cuda", generator=g)
    gr = torch.randn(n, device="cuda", generator=g) * 64
    buf = torch.randn(n, device="cuda", generator=g)
    inv = torch.tensor(1 / 64.0, device="cuda")
    rp, rb = p.clone(), buf.clone()
    _flat_sgd(rp, gr, rb, lr=0.1, momentum=0.9, dampening=0.0, weight_decay=1e-4, nesterov=nesterov,
              maximize=False, first_step=first, inv_scale=inv)
    shadow = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    g2 = gr.clone()
    ops.sgd_step(p, g2, buf, shadow, lr=0.1, momentum=0.9, dampening=0.0, weight_decay=1e-4, nesterov=nesterov,
                 maximize=False, first_step=first, inv_scale=inv, found_inf=torch.zeros((), device="cuda"),
                 zero_grad=True)
    assert torch.allclose(p, rp, rtol=1e-5, atol=1e-6) and torch.allclose(buf, rb, rtol=1e-5, atol=1e-6)
    assert torch.equal(shadow, p.to(torch.bfloat16)) and float(g2.abs().max()) == 0.0


def test_sgd_skips_on_found_inf_and_check_finite():
    n = 4096
    p = torch.randn(n, device="cuda")
    g = torch.randn(n, device="cuda")
    buf = torch.randn(n, device="cuda")
    g[77] = float("inf")
    flag = torch.zeros((), device="cuda")
    ops.check_finite(g, flag)
    assert float(flag) == 1.0
    p0, b0 = p.clone(), buf.clone()
    ops.sgd_step(p, g, buf, None, lr=0.1, momentum=0.9, dampening=0.0, weight_decay=0.0, nesterov=False,
                 maximize=False, first_step=False, found_inf=flag)
    assert torch.equal(p, p0) and torch.equal(buf, b0)
    ok = torch.zeros((), device="cuda")
    ops.check_finite(torch.randn(n, device="cuda"), ok)
    assert float(ok) == 0.0


@pytest.mark.parametrize("dt", DT)
def test_xent_and_topk(dt):
    g = torch.Generator(device="cuda").manual_seed(1)
    logits = (torch.randn(400, 1000, device="cuda", generator=g) * 3).to(dt)
    labels = torch.randint(0, 1000, (400,), device="cuda", generator=g)
    scale = torch.tensor(128.0, device="cuda")
    loss, dl = ops.xent_fwd_bwd(logits, labels, grad_scale=scale)
    ref_in = logits.float().requires_grad_(True)
    ref = F.cross_entropy(ref_in, labels)
    ref.backward()
    assert abs(float(loss) - float(ref)) < 1e-3 * abs(float(ref))
    assert rel_err(dl, ref_in.grad * 128.0) < (1e-2 if dt == torch.bfloat16 else 2e-3)
    hits = ops.topk_correct(logits, labels, (1, 5))
    _, pred = logits.float().topk(5, -1, True, True)
    eq = pred.eq(labels[:, None])
    assert hits.tolist() == [float(eq[:, :1].sum()), float(eq.sum())]


def test_fused_xent_survives_the_initial_amp_scale():
    """fp16 logits, GradScaler's initial scale 65536 (does not fit fp16) and batch 400: the gradient must come out
    finite and un-flushed — the loss scale is applied in fp32, before the rounding to 16 bits."""
    from pytorch_distributed_b200.nn import CrossEntropyLoss
    g = torch.Generator(device="cuda").manual_seed(11)
    logits = (torch.randn(400, 1000, device="cuda", generator=g) * 4).half().requires_grad_(True)
    labels = torch.randint(0, 1000, (400,), device="cuda", generator=g)
    scale = torch.tensor(65536.0, device="cuda")
    (CrossEntropyLoss()(logits, labels) * scale).backward()
    ref_in = logits.detach().float().requires_grad_(True)
    (F.cross_entropy(ref_in, labels) * 65536.0).backward()
    assert logits.grad.dtype == torch.float16 and bool(torch.isfinite(logits.grad).all())
    assert rel_err(logits.grad, ref_in.grad) < 2e-3
    # entries that (softmax - onehot)/400 alone would flush to zero in fp16 (< 6e-8) survive with the scale applied first
    tiny = (ref_in.grad.abs() / 65536.0 < 5.9e-8) & (ref_in.grad.abs() > 6.2e-5)
    assert int(tiny.sum()) > 0 and bool((logits.grad[tiny] != 0).all())


def test_fused_xent_takes_fp32_logits():
    from pytorch_distributed_b200.nn import CrossEntropyLoss
    g = torch.Generator(device="cuda").manual_seed(12)
    logits = (torch.randn(64, 1000, device="cuda", generator=g) * 3).requires_grad_(True)
    labels = torch.randint(0, 1000, (64,), device="cuda", generator=g)
    loss = CrossEntropyLoss()(logits, labels)
    loss.backward()
    ref_in = logits.detach().clone().requires_grad_(True)
    ref = F.cross_entropy(ref_in, labels)
    ref.backward()
    assert abs(float(loss) - float(ref)) < 1e-5 * abs(float(ref)) + 1e-6
    assert rel_err(logits.grad, ref_in.grad) < 1e-5


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("rows,C", [(4 * 56 * 56, 64), (3 * 14 * 14, 1024), (5 * 49, 2048), (1000, 24)])
def test_bn_forward(dt, rows, C):
    g = torch.Generator(device="cuda").manual_seed(2)
    x = (torch.randn(rows, C, device="cuda", generator=g) * 2 + 0.5).to(dt)
    res = torch.randn(rows, C, device="cuda", generator=g).to(dt)
    gamma = torch.rand(C, device="cuda", generator=g) + 0.5
    beta = torch.randn(C, device="cuda", generator=g)
    sums = torch.empty(2 * C, device="cuda")
    ops.bn_stats(x, sums)
    ops.bn_stats(x, sums)                       # scratch ticket must have been reset by the first call
    xf = x.float()
    assert rel_err(sums[:C], xf.sum(0)) < 1e-4 and rel_err(sums[C:], (xf * xf).sum(0)) < 1e-4
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    nbt = torch.zeros((), dtype=torch.int64, device="cuda")
    mean, rstd = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
    y = torch.empty_like(x)
    pow2 = (C & (C - 1)) == 0
    bits = torch.zeros(rows * C // 8, dtype=torch.uint8, device="cuda") if pow2 else None
    ops.bn_apply(x, res, y, sums, rows, gamma, beta, 1e-5, 0.1, rm, rv, nbt, mean, rstd, True, relu_mask=bits)
    if pow2:       # bit k of byte i <-> element 8*i + k passed the ReLU
        want = ((y.reshape(-1, 8) > 0).to(torch.int32) << torch.arange(8, device="cuda", dtype=torch.int32)).sum(1)
        assert torch.equal(bits.to(torch.int32), want)
    rrm, rrv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    ref = F.batch_norm(xf, rrm, rrv, gamma, beta, True, 0.1, 1e-5)
    ref = F.relu(ref + res.float())
    assert rel_err(y, ref) < (1.5e-2 if dt == torch.bfloat16 else 2e-3)
    assert torch.allclose(rm, rrm, rtol=1e-3, atol=1e-4) and torch.allclose(rv, rrv, rtol=1e-3, atol=1e-4)
    assert int(nbt) == 1
    assert torch.allclose(mean, xf.mean(0), rtol=1e-3, atol=1e-4)
    ye = torch.empty_like(x)
    ops.bn_apply_eval(x, None, ye, rm, rv, gamma, beta, 1e-5, False)
    assert rel_err(ye, F.batch_norm(xf, rm, rv, gamma, beta, False, 0.0, 1e-5)) < 1.5e-2


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("rows,C,with_z", [(4 * 28 * 28, 128, True), (2 * 49, 2048, True), (3000, 256, False)])
def test_bn_backward(dt, rows, C, with_z):
    g = torch.Generator(device="cuda").manual_seed(3)
    x = (torch.randn(rows, C, device="cuda", generator=g) * 1.5 + 0.3).to(dt)
    gamma = torch.rand(C, device="cuda", generator=g) + 0.5
    beta = torch.randn(C, device="cuda", generator=g) * 0.1
    dz = torch.randn(rows, C, device="cuda", generator=g).to(dt)
    xr = x.float().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    pre = F.batch_norm(xr, None, None, gr, br, True, 0.1, 1e-5)
    z = F.relu(pre) if with_z else pre
    z.backward(dz.float())
    xf = x.float()
    mean, var = xf.mean(0), xf.var(0, unbiased=False)
    rstd = torch.rsqrt(var + 1e-5)
    zt = z.detach().to(dt) if with_z else None
    dsums = torch.zeros(2 * C, device="cuda")
    ops.bn_bwd_reduce(dz, zt, x, mean, rstd, dsums)
    dx, dres = torch.empty_like(x), torch.empty_like(x)
    dgamma, dbeta = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    ops.bn_bwd_apply(dz, zt, x, dx, dres, mean, rstd, gamma, dsums, rows, dgamma, dbeta)
    if with_z:     # the bit-mask variant must agree with the z-based one
        bits = ((zt.reshape(-1, 8) > 0).to(torch.int32) << torch.arange(8, device="cuda", dtype=torch.int32)).sum(1).to(torch.uint8)
        ds2 = torch.zeros(2 * C, device="cuda")
        ops.bn_bwd_reduce(dz, None, x, mean, rstd, ds2, relu_mask=bits)
        assert rel_err(ds2, dsums) < 1e-4
        # replicated accumulators + last-block fold (the contention-free path the engine uses)
        ds3 = torch.full((2 * C,), float("nan"), device="cuda")
        scratch = torch.zeros(ops.bn_bwd_reduce_scratch_floats(C), device="cuda")
        ops.bn_bwd_reduce(dz, None, x, mean, rstd, ds3, relu_mask=bits, scratch=scratch)
        assert rel_err(ds3, dsums) < 1e-4
        dx2, dres2 = torch.empty_like(x), torch.empty_like(x)
        ops.bn_bwd_apply(dz, None, x, dx2, dres2, mean, rstd, gamma, dsums, rows, None, None, relu_mask=bits)
        assert torch.equal(dx2, dx) and torch.equal(dres2, dres)
    tol = 2e-2 if dt == torch.bfloat16 else 4e-3
    assert rel_err(dgamma, gr.grad) < tol and rel_err(dbeta, br.grad) < tol
    assert rel_err(dx, xr.grad) < tol
    mask = (z.detach() > 0).float() if with_z else 1.0
    assert rel_err(dres, dz.float() * mask) < 1e-6


@pytest.mark.parametrize("dt", DT)
def test_pools(dt):
    g = torch.Generator(device="cuda").manual_seed(4)
    x = torch.randn(3, 20, 20, 64, device="cuda", generator=g).to(dt)
    x[0, :4, :4] = 0                                          # ties (post-ReLU zeros)
    y = torch.empty(3, 10, 10, 64, device="cuda", dtype=dt)
    am = torch.empty(3, 10, 10, 64, device="cuda", dtype=torch.uint8)
    ops.maxpool_fwd(x, y, am)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    ref = F.max_pool2d(xr, 3, 2, 1)
    assert torch.equal(y.float(), ref.permute(0, 2, 3, 1))
    dy = torch.randn(3, 10, 10, 64, device="cuda", generator=g).to(dt)
    ref.backward(dy.float().permute(0, 3, 1, 2))
    dx = torch.empty_like(x)
    ops.maxpool_bwd(dy, am, dx)
    assert rel_err(dx, xr.grad.permute(0, 2, 3, 1)) < 1e-2
    a = torch.randn(5, 7, 7, 128, device="cuda", generator=g).to(dt)
    o = torch.empty(5, 128, device="cuda", dtype=dt)
    ops.avgpool_fwd(a, o)
    assert rel_err(o, a.float().mean((1, 2))) < 1e-2
    do = torch.randn(5, 128, device="cuda", generator=g).to(dt)
    da = torch.empty_like(a)
    ops.avgpool_bwd(do, da)
    assert rel_err(da, (do.float() / 49)[:, None, None, :].expand(5, 7, 7, 128)) < 1e-2


@pytest.mark.parametrize("dt", DT)
def test_layout_helpers(dt):
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(2, 3, 32, 32, device="cuda", generator=g)
    y = torch.empty(2, 32, 32, 3, device="cuda", dtype=dt)
    ops.nchw_to_nhwc16(x, y)
    assert torch.equal(y, x.permute(0, 2, 3, 1).to(dt))
    col = torch.empty(2 * 16 * 16, 152, device="cuda", dtype=dt)
    ops.stem_im2col(x, col)
    ref = F.unfold(x, 7, padding=3, stride=2)                  # [N, C*49, L], rows ordered (c, kh, kw)
    ref = ref.view(2, 3, 49, 256).permute(0, 3, 2, 1).reshape(512, 147)   # -> (kh*7+kw)*3 + c
    assert torch.equal(col[:, :147], ref.to(dt)) and float(col[:, 147:].abs().max()) == 0.0
    a = torch.randn(2, 8, 8, 16, device="cuda", generator=g).to(dt)
    b = torch.randn(2, 8, 8, 16, device="cuda", generator=g).to(dt)
    o = torch.empty_like(a)
    ops.add16(a, b, o)
    assert torch.equal(o, (a.float() + b.float()).to(dt))
    s = torch.randn(2, 4, 4, 16, device="cuda", generator=g).to(dt)
    up = torch.empty(2, 8, 8, 16, device="cuda", dtype=dt)
    ops.upsample2_zero(s, up)
    ref_up = torch.zeros(2, 8, 8, 16, device="cuda", dtype=dt)
    ref_up[:, ::2, ::2] = s
    assert torch.equal(up, ref_up)
    ops.add_strided2(a, s, o)
    assert torch.equal(o, (a.float() + ref_up.float()).to(dt))
    w = torch.randn(8, 3, 3, 16, device="cuda", generator=g).to(dt)
    wt = torch.empty(16, 3, 3, 8, device="cuda", dtype=dt)
    ops.weight_dgrad_layout(w, wt, 8, 3, 3, 16)
    assert torch.equal(wt, w.flip(1, 2).permute(3, 1, 2, 0).contiguous())
    pd = torch.empty(8, 152, device="cuda", dtype=dt)
    ops.pad_rows16(w.reshape(8, 144), pd, 8, 144, 152)
    assert torch.equal(pd[:, :144], w.reshape(8, 144)) and float(pd[:, 144:].abs().max()) == 0
    cs = torch.zeros(16, device="cuda")
    ops.colsum16(a.reshape(-1, 16), cs)
    assert rel_err(cs, a.float().reshape(-1, 16).sum(0)) < 1e-5


def test_scaler_update_kernel_matches_tensor_path():
    """The one-thread amp scale update (K13) against the tensor-select implementation the CPU path uses."""
    from pytorch_distributed_b200 import amp
    dev = torch.device("cuda")
    events = [0, 0, 1, 0, 0, 0, 1, 1, 0, 0, 0, 0, 0]          # found_inf per step
    a = amp.GradScaler(init_scale=1024.0, growth_interval=3, device=dev)
    b = amp.GradScaler(init_scale=1024.0, growth_interval=3, device=torch.device("cpu"))
    a._lazy_init(dev), b._lazy_init(torch.device("cpu"))
    for inf in events:
        a._found_inf.fill_(float(inf)), b._found_inf.fill_(float(inf))
        a.update(), b.update()
        assert a.get_scale() == b.get_scale()
        assert int(a._growth_tracker.item()) == int(b._growth_tracker.item())
        assert float(a._found_inf.item()) == 0.0
        assert float(a._inv_scale.item()) == pytest.approx(1.0 / a.get_scale(), rel=1e-7)
    # growth saturates instead of overflowing
    c = amp.GradScaler(init_scale=3.0e38, growth_interval=1, device=dev)
    c._lazy_init(dev)
    c.update()
    assert c.get_scale() == pytest.approx(3.0e38)


def test_weight_dgrad_layout_batched_matches_permute():
    """wt[c][r][s][k] = w[k][R-1-r][S-1-s][c] for a table of layers in one launch (tiled transpose), incl. sizes that
    are not multiples of the 32 x 32 tile."""
    layers = [(40, 3, 3, 24), (64, 1, 1, 64), (1000, 1, 1, 2048), (128, 3, 3, 128), (8, 7, 7, 8)]
    g = torch.Generator(device="cuda").manual_seed(3)
    n = sum(K * R * S * C for K, R, S, C in layers)
    w = torch.randn(n, device="cuda", generator=g).bfloat16()
    wt = torch.full((n,), float("nan"), device="cuda", dtype=torch.bfloat16)
    rows, off = [], 0
    for K, R, S, C in layers:
        rows.append([off, off, K, R, S, C])
        off += K * R * S * C
    table = torch.tensor(rows, dtype=torch.int64, device="cuda")
    ops.weight_dgrad_layout_batched(w, wt, table, len(layers))
    torch.cuda.synchronize()
    for o, _, K, R, S, C in rows:
        src = w[o:o + K * R * S * C].view(K, R, S, C)
        want = src.flip(1, 2).permute(3, 1, 2, 0).contiguous()          # [C, R, S, K]
        assert torch.equal(wt[o:o + K * R * S * C].view(C, R, S, K), want), (K, R, S, C)


def test_fp16_subnormals_survive_bn_backward():
    """The library is compiled with -ftz=true (fp32 denormals flush to zero).  That must not touch fp16 SUBNORMAL
    gradients (|g| < 6.1e-5, what loss scaling exists to protect): widened to fp32 they are ordinary normal numbers, and
    the residual-branch gradient the BN backward apply kernel passes through (dres = dz under the ReLU mask) must come
    back bit for bit."""
    rows, C = 4096, 64
    g = torch.Generator(device="cuda").manual_seed(21)
    x = torch.randn(rows, C, device="cuda", generator=g).half()
    tiny = (torch.rand(rows, C, device="cuda", generator=g) * 5.9e-5 + 6e-8).half()       # all below the fp16 normal range
    assert float(tiny.float().abs().max()) < 6.104e-5 and float(tiny.float().abs().min()) > 0
    bits = torch.full((rows * C // 8,), 255, dtype=torch.uint8, device="cuda")            # ReLU passed everywhere
    mean, rstd, gamma = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda"), torch.ones(C, device="cuda")
    ds = torch.zeros(2 * C, device="cuda")
    dx, dres = torch.empty_like(x), torch.empty_like(x)
    ops.bn_bwd_apply(tiny, None, x, dx, dres, mean, rstd, gamma, ds, rows, None, None, relu_mask=bits)
    torch.cuda.synchronize()
    assert torch.equal(dres, tiny)
    assert torch.equal(dx, tiny)            # zero sums, gamma = rstd = 1: dx = g exactly
