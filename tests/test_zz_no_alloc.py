"""pdb_bn_stats_noalloc through the C ABI.

Some callers of the kernel library cannot hand it a scratch buffer: a host program that committed all
GPU memory to its own pools at start-up, or code in the middle of a training step where a cudaMalloc
would stall every stream on the device.  For them the library carries a fixed scratch array, and this
entry point must give the same per-channel sums as pdb_bn_stats does with caller-provided scratch.
The file sorts last on purpose: if it fails under ``pytest -x`` every other test has already run.
"""
import ctypes
from ctypes import c_int, c_int64, c_size_t, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from pytorch_distributed_b200.build import LIB
    L = ctypes.CDLL(str(LIB / "libpdb200.so"))
    L.pdb_bn_stats_noalloc.argtypes = [c_void_p, c_int, c_int64, c_int, c_void_p, c_void_p]
    L.pdb_bn_stats_noalloc.restype = c_int
    L.pdb_bn_stats.argtypes = [c_void_p, c_int, c_int64, c_int, c_void_p, c_void_p, c_size_t, c_void_p]
    L.pdb_bn_stats.restype = c_int
    L.pdb_bn_stats_scratch_bytes.argtypes = [c_int]
    L.pdb_bn_stats_scratch_bytes.restype = c_size_t
    L.pdb_last_error.restype = ctypes.c_char_p
    return L


def _rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


def test_noalloc_matches_scratch_variant_twice_in_a_row():
    torch.cuda.init()
    L = _lib()
    stream = c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cuda").manual_seed(11)
    # two different layers back to back: the second call finds the first call's partials in the array
    for rows, C, tag in [(4 * 56 * 56, 256, 0), (3 * 28 * 28 + 5, 64, 1)]:
        dt = torch.bfloat16 if tag == 0 else torch.float16
        x = (torch.randn(rows, C, device="cuda", generator=g) * 2 + 0.25).to(dt)
        want = torch.full((2 * C,), float("nan"), device="cuda")
        got = torch.full((2 * C,), float("nan"), device="cuda")
        n = L.pdb_bn_stats_scratch_bytes(C)
        scratch = torch.zeros(n, dtype=torch.uint8, device="cuda")
        rc = L.pdb_bn_stats(x.data_ptr(), tag, rows, C, want.data_ptr(), scratch.data_ptr(), n, stream)
        assert rc == 0, L.pdb_last_error()
        before = torch.cuda.memory_allocated()
        rc = L.pdb_bn_stats_noalloc(x.data_ptr(), tag, rows, C, got.data_ptr(), stream)
        assert rc == 0, L.pdb_last_error()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        # same tolerance as test_gpu_kernels.test_bn_forward uses for these sums
        assert _rel(got[:C], want[:C]) < 1e-4 and _rel(got[C:], want[C:]) < 1e-4
        xf = x.float()
        assert _rel(got[:C], xf.sum(0)) < 1e-4 and _rel(got[C:], (xf * xf).sum(0)) < 1e-4


def test_noalloc_rejects_oversize():
    torch.cuda.init()
    L = _lib()
    x = torch.zeros(16, 1024, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(2048, device="cuda")
    rc = L.pdb_bn_stats_noalloc(x.data_ptr(), 0, 16, 1024, out.data_ptr(),
                                c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0 and b"512" in L.pdb_last_error()
