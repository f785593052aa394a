"""fp32 PyTorch oracles for the sm_100a kernels + small runners shared by tests and probes."""
import torch
import torch.nn.functional as F

from pytorch_distributed_b200 import ops
from pytorch_distributed_b200.ops import ConvShape


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.float(), b.float()
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


def conv_case(N, H, W, C, K, R, S, stride, pad, dtype=torch.bfloat16, block_n=0, stats=True, seed=0):
    """Runs pdb_conv_fprop and returns (rel_err_y, rel_err_stats)."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(N, H, W, C, device="cuda", generator=g).to(dtype)
    w = (torch.randn(K, R, S, C, device="cuda", generator=g) / (R * S * C) ** 0.5).to(dtype)
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    y = torch.full((N, P, Q, K), float("nan"), device="cuda", dtype=dtype)
    st = torch.zeros(2 * K, device="cuda") if stats else None
    ops.conv_fprop(x, w, y, ConvShape(N, H, W, C, K, R, S, stride, pad), stats=st, block_n=block_n)
    torch.cuda.synchronize()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), stride=stride, padding=pad)
    ref = ref.permute(0, 2, 3, 1)
    e_y = rel_err(y, ref)
    e_s = 0.0
    if stats:
        yf = y.float().reshape(-1, K)
        e_s = max(rel_err(st[:K], yf.sum(0)), rel_err(st[K:], (yf * yf).sum(0)))
        # the contention-free variant the engine uses: replicated accumulators folded by the last CTA
        st2 = torch.full((2 * K,), float("nan"), device="cuda")
        scratch = torch.zeros(16 * 2 * K + 32, device="cuda")
        ops.conv_fprop(x, w, y, ConvShape(N, H, W, C, K, R, S, stride, pad), stats=st2, stats_scratch=scratch,
                       block_n=block_n)
        torch.cuda.synchronize()
        e_s = max(e_s, rel_err(st2[:K], yf.sum(0)), rel_err(st2[K:], (yf * yf).sum(0)))
    return e_y, e_s


def wgrad_case(N, H, W, C, K, R, S, stride, pad, dtype=torch.bfloat16, block_n=0, splits=0, seed=0):
    torch.backends.cudnn.allow_tf32 = False
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(N, H, W, C, device="cuda", generator=g).to(dtype)
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    dy = torch.randn(N, P, Q, K, device="cuda", generator=g).to(dtype)
    dw = torch.zeros(K, R, S, C, device="cuda")
    ops.conv_wgrad(x, dy, dw, ConvShape(N, H, W, C, K, R, S, stride, pad), splits=splits, block_n=block_n)
    torch.cuda.synchronize()
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(False)
    wr = torch.zeros(K, C, R, S, device="cuda", requires_grad=True)
    out = F.conv2d(xr, wr, stride=stride, padding=pad)
    out.backward(dy.float().permute(0, 3, 1, 2))
    return rel_err(dw, wr.grad.permute(0, 2, 3, 1))


def dgrad_case(N, H, W, C, K, R, S, pad, dtype=torch.bfloat16, seed=0):
    """stride-1 dgrad through conv_fprop with the flipped/transposed weights."""
    torch.backends.cudnn.allow_tf32 = False
    g = torch.Generator(device="cuda").manual_seed(seed)
    w = (torch.randn(K, R, S, C, device="cuda", generator=g) / (R * S * C) ** 0.5).to(dtype)
    P, Q = H + 2 * pad - R + 1, W + 2 * pad - S + 1
    dy = torch.randn(N, P, Q, K, device="cuda", generator=g).to(dtype)
    wt = torch.empty(C, R, S, K, device="cuda", dtype=dtype)
    ops.weight_dgrad_layout(w, wt, K, R, S, C)
    dx = torch.full((N, H, W, C), float("nan"), device="cuda", dtype=dtype)
    ops.conv_fprop(dy, wt, dx, ConvShape(N, P, Q, K, C, R, S, 1, R - 1 - pad))
    torch.cuda.synchronize()
    xr = torch.zeros(N, C, H, W, device="cuda", requires_grad=True)
    F.conv2d(xr, w.float().permute(0, 3, 1, 2), padding=pad).backward(dy.float().permute(0, 3, 1, 2))
    return rel_err(dx, xr.grad.permute(0, 2, 3, 1))


def stem_case(N=4, H=64, W=64, dtype=torch.bfloat16, seed=0):
    """The 7x7/2 stem through the space-to-depth TMA path vs fp32 F.conv2d: (rel err y, stats, dW)."""
    torch.backends.cudnn.allow_tf32 = False
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(N, 3, H, W, device="cuda", generator=g)
    w = (torch.randn(64, 7, 7, 3, device="cuda", generator=g) / 147 ** 0.5)
    w16 = w.to(dtype)
    Hs, Ws = H // 2, W // 2
    xs = torch.empty(N, Hs, Ws, 16, device="cuda", dtype=dtype)
    ws = torch.empty(64, 256, device="cuda", dtype=dtype)
    y = torch.full((N, Hs, Ws, 64), float("nan"), device="cuda", dtype=dtype)
    st = torch.full((128,), float("nan"), device="cuda")
    scratch = torch.zeros(16 * 128 + 32, device="cuda")
    ops.stem_s2d(x, xs)
    ops.stem_weight_s2d(w16, ws)
    ops.stem_fprop(xs, ws, y, stats=st, stats_scratch=scratch)
    torch.cuda.synchronize()
    xr = x.to(dtype).float()
    wr = w16.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)      # [64, 3, 7, 7]
    ref = F.conv2d(xr, wr, stride=2, padding=3)
    e_y = rel_err(y, ref.detach().permute(0, 2, 3, 1))
    yf = y.float().reshape(-1, 64)
    e_s = max(rel_err(st[:64], yf.sum(0)), rel_err(st[64:], (yf * yf).sum(0)))
    dy = torch.randn(N, Hs, Ws, 64, device="cuda", generator=g).to(dtype)
    dws = torch.zeros(64, 256, device="cuda")
    dw = torch.full((64, 7, 7, 3), 1.0, device="cuda")
    ops.stem_wgrad(xs, dy, dws)
    ops.stem_dw_unpack(dws, dw, accumulate=True)
    torch.cuda.synchronize()
    ref.backward(dy.float().permute(0, 3, 1, 2))
    e_w = rel_err(dw - 1.0, wr.grad.permute(0, 2, 3, 1))
    return e_y, e_s, e_w
