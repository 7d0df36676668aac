"""tcgen05 operand-convention self-test (single-CTA GEMM through the C ABI) vs torch fp32 matmul."""
import ctypes as C

import pytest
import torch

from conftest import pkg

pytestmark = pytest.mark.gpu


def _run(mode, N, K):
    _lib = pkg("_lib")
    lib = _lib.load()
    g = torch.Generator().manual_seed(mode * 1000 + N + K)
    A = torch.randn(128, K, generator=g)
    B = torch.randn(N, K, generator=g)
    # make every element distinguishable so a layout error cannot hide
    A += torch.arange(128).float()[:, None] * 0.01 + torch.arange(K).float()[None] * 0.1
    ref = A.double() @ B.double().T
    Ad, Bd = A.cuda(), B.cuda()
    if mode == 3:
        ref = A.double().sum(1, keepdim=True).expand(128, N)
    if mode >= 2:
        Ad, Bd = Ad.t().contiguous(), Bd.t().contiguous()  # [K,128], [K,N]
    D = torch.full((128, N), float("nan"), device="cuda")
    _lib.check(lib.d3g_selftest_umma(mode, N, K, _lib.ptr(Ad), _lib.ptr(Bd), _lib.ptr(D),
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    err = (D.cpu().double() - ref).abs().max().item() / ref.abs().max().item()
    return err


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("N,K", [(256, 64), (64, 32), (32, 32)])
def test_umma_selftest(mode, N, K):
    err = _run(mode, N, K)
    assert err < 2e-3, "mode %d N=%d K=%d rel err %.3e" % (mode, N, K, err)


@pytest.mark.parametrize("N,K", [(96, 40), (224, 64)])
def test_umma_mn_major_odd_shapes(N, K):
    assert _run(2, N, K) < 2e-3


@pytest.mark.parametrize("K", [8, 32, 64])
def test_umma_ones_column_sum(K):
    assert _run(3, 16, K) < 2e-3


def test_umma_k8_steps():
    # K not a multiple of the 32-wide panel: exercises the +32 B descriptor advance inside a swizzle atom
    for mode in (0, 2):
        assert _run(mode, 128, 8) < 2e-3
        assert _run(mode, 128, 24) < 2e-3


def _run_f16(mode, N, K):
    _lib = pkg("_lib")
    lib = _lib.load()
    g = torch.Generator().manual_seed(mode * 1000 + N + K)
    # fp16-representable values, every element distinguishable
    A = (torch.randn(128, K, generator=g) + torch.arange(128).float()[:, None] * 0.01
         + torch.arange(K).float()[None] * 0.1).half().float()
    B = torch.randn(N, K, generator=g).half().float()
    ref = A.double() @ B.double().T
    Ad, Bd = A.cuda(), B.cuda()
    if mode == 12:
        Ad, Bd = Ad.t().contiguous(), Bd.t().contiguous()  # [K,128], [K,N]
    D = torch.full((128, N), float("nan"), device="cuda")
    _lib.check(lib.d3g_selftest_umma_f16(mode, N, K, _lib.ptr(Ad), _lib.ptr(Bd), _lib.ptr(D),
                                         C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return (D.cpu().double() - ref).abs().max().item() / ref.abs().max().item()


@pytest.mark.parametrize("mode", [10, 11, 12])
@pytest.mark.parametrize("N,K", [(256, 64), (64, 32), (128, 128)])
def test_umma_f16_selftest(mode, N, K):
    """kind::f16 with exact fp16 operands and fp32 accumulation: only accumulation rounding remains."""
    err = _run_f16(mode, N, K)
    assert err < 1e-5, "mode %d N=%d K=%d rel err %.3e" % (mode, N, K, err)


def test_umma_f16_k16_steps():
    for mode in (10, 12):
        assert _run_f16(mode, 64, 16) < 1e-5
        assert _run_f16(mode, 64, 48) < 1e-5
    assert _run_f16(10, 16, 64) < 1e-5
