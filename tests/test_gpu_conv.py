"""tcgen05 implicit-GEMM convolutions vs fp32 F.conv2d on the ResNet-50 shape families."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from ._gpu_helpers import conv_case, dgrad_case, wgrad_case  # noqa: E402

TOL = 1.2e-2       # 16-bit output rounding (2^-8 relative) on top of fp32 accumulation


@pytest.mark.parametrize("case", [
    (2, 16, 16, 64, 64, 1, 1, 1, 0, 64), (2, 16, 16, 128, 256, 1, 1, 1, 0, 256),
    (3, 14, 14, 256, 1024, 1, 1, 1, 0, 256), (2, 56, 56, 64, 64, 3, 3, 1, 1, 64),
    (2, 28, 28, 128, 128, 3, 3, 1, 1, 128), (2, 56, 56, 128, 128, 3, 3, 2, 1, 128),
    (2, 56, 56, 256, 512, 1, 1, 2, 0, 256), (5, 7, 7, 512, 512, 3, 3, 1, 1, 256),
    (400, 1, 1, 2048, 1000, 1, 1, 1, 0, 128), (2, 112, 112, 152, 64, 1, 1, 1, 0, 64),
])
def test_fprop(case):
    N, H, W, C, K, R, S, stride, pad, bn = case
    e_y, e_s = conv_case(N, H, W, C, K, R, S, stride, pad, block_n=bn)
    assert e_y < TOL and e_s < 1e-3, (e_y, e_s)


def test_fprop_fp16():
    e_y, e_s = conv_case(2, 28, 28, 128, 128, 3, 3, 1, 1, dtype=torch.float16)
    assert e_y < 3e-3 and e_s < 1e-3


@pytest.mark.parametrize("case", [(2, 28, 28, 128, 128, 3, 3, 1), (2, 14, 14, 256, 1024, 1, 1, 0)])
def test_dgrad(case):
    assert dgrad_case(*case) < TOL


@pytest.mark.parametrize("case", [
    (2, 16, 16, 64, 64, 1, 1, 1, 0, 64), (2, 28, 28, 64, 64, 3, 3, 1, 1, 256),
    (2, 28, 28, 128, 256, 3, 3, 1, 1, 256), (2, 56, 56, 128, 128, 3, 3, 2, 1, 256),
    (2, 56, 56, 256, 512, 1, 1, 2, 0, 256), (400, 1, 1, 2048, 1000, 1, 1, 1, 0, 256),
])
def test_wgrad(case):
    N, H, W, C, K, R, S, stride, pad, bn = case
    assert wgrad_case(N, H, W, C, K, R, S, stride, pad, block_n=bn) < 5e-3
