"""tcgen05 GEMM kernels vs a plain fp32/fp64 PyTorch reference."""
import pytest
import torch

from omldm_b200.ops import gemm

pytestmark = pytest.mark.gpu


def _tf32_ok(out, a, b_t, extra=1.0):
    """The tensor core truncates fp32 operands to tf32 (10 mantissa bits): each product is off by at
    most ~2^-9 relative, so |err_ij| <= 2^-9 * (|a| |b|)_ij plus fp32 accumulation noise."""
    ref = a.double() @ b_t.double()
    bound = (a.double().abs() @ b_t.double().abs()) * (2.0 ** -9) * extra + 1e-5
    return bool(((out.double() - ref).abs() <= bound).all())


@pytest.mark.parametrize("M,N,K", [(128, 32, 32), (128, 256, 64), (100, 10, 1000), (300, 130, 257),
                                   (4096, 16, 4096), (257, 256, 100000)])
def test_gemm_tn(M, N, K):
    torch.manual_seed(0)
    Kp = (K + 3) // 4 * 4
    a = torch.randn(M, Kp, device="cuda")[:, :K]
    b = torch.randn(N, Kp, device="cuda")[:, :K]
    out = gemm.gemm_tn(a, b)
    assert _tf32_ok(out, a, b.t())


@pytest.mark.parametrize("M,N,K,ksplit", [(128, 32, 32, 1), (128, 256, 64, 1), (1024, 1024, 512, 1),
                                          (1002, 1002, 1000, 3), (434, 434, 777, 2), (32, 64, 4096, 4)])
def test_gemm_outer(M, N, K, ksplit):
    torch.manual_seed(1)
    lda = (M + 31) // 32 * 32
    ldb = (N + 31) // 32 * 32
    a = torch.randn(K, lda, device="cuda")
    b = torch.randn(K, ldb, device="cuda")
    out = gemm.gemm_outer(a, b, m=M, n=N, ksplit=ksplit)
    assert _tf32_ok(out, a[:, :M].t(), b[:, :N])


def test_gemm_outer_accumulate_upper():
    torch.manual_seed(2)
    K, d = 2048, 1024
    z = torch.randn(K, d, device="cuda")
    g = torch.zeros(d, d, device="cuda")
    for _ in range(3):
        gemm.gemm_outer(z, z, g, accumulate=True, ksplit=4, upper_only=True)
    ref = 3 * (z.double().t() @ z.double())
    bound = 3 * (z.double().abs().t() @ z.double().abs()) * 2.0 ** -9 + 1e-4
    iu = torch.triu(torch.ones(d, d, device="cuda", dtype=torch.bool))
    assert bool((((g.double() - ref).abs() <= bound) | ~iu).all())
