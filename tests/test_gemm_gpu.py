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


@pytest.mark.parametrize("M,N,K", [(256, 20, 16), (300, 17, 8), (1000, 130, 64), (128, 256, 300), (4096, 33, 257)])
def test_gemm_nn_mixed_majors(M, N, K):
    """A K-major, B MN-major (both row-major in memory): the back-propagation GEMM of the NN learner."""
    torch.manual_seed(3)
    lda = (K + 3) // 4 * 4
    ldb = (N + 31) // 32 * 32
    a = torch.randn(M, lda, device="cuda")[:, :K]
    b = torch.randn(K, ldb, device="cuda")
    out = torch.full((M, ldb), float("nan"), device="cuda")[:, :N]
    gemm.gemm_nn(a, b, out)
    assert _tf32_ok(out, a, b[:, :N])
    gemm.gemm_nn(a, b, out, accumulate=True)
    assert _tf32_ok(out, a, 2 * b[:, :N])


@pytest.mark.parametrize("B,d,K", [(2048, 32, 16), (1000, 128, 256), (16384, 128, 256), (300, 20, 100), (129, 7, 33)])
def test_kmeans_assignment_in_the_gemm_epilogue(B, d, K):
    """Distance GEMM + argmin from TMEM + one-hot rows in one kernel against plain PyTorch: same assignment wherever
    the two nearest centroids are not within tf32 rounding of each other, same inertia, exact one-hot rows."""
    from omldm_b200.ops import functional as F
    torch.manual_seed(B + K)
    ld = F.pad32(d + 2)
    Z = torch.zeros(B, ld, device="cuda")
    cent = torch.randn(K, d, device="cuda") * 3
    lab = torch.randint(0, K, (B,), device="cuda")
    Z[:, :d] = cent[lab] + 0.3 * torch.randn(B, d, device="cuda")
    Z[:, d] = 1.0
    Z[::11, :] = 0.0                                   # empty rows
    Cm = torch.zeros(K, ld, device="cuda")
    Cm[:, :d] = cent
    cnorm = (cent ** 2).sum(1)
    cnorm[K - 1] = float("inf")                        # an empty centroid is never chosen
    T = torch.full((B, F.pad32(K)), float("nan"), device="cuda")
    assign = torch.full((B,), -7, dtype=torch.int32, device="cuda")
    stats = torch.zeros(4, dtype=torch.float64, device="cuda")
    assert F.kmeans_assign_fused(Z, d, Cm, cnorm, K, T, stats, assign)
    torch.cuda.synchronize()
    valid = Z[:, d] != 0
    D2 = cnorm[None, :].double() - 2.0 * (Z[:, :d].double() @ cent.double().t())
    ref = D2.argmin(1)
    srt = D2.sort(1).values
    clear = (srt[:, 1] - srt[:, 0]) > 1e-2 * (Z[:, :d].double().abs() @ cent.double().abs().t()).max(1).values * 2.0 ** -9 * 4
    a = assign.long()
    assert bool((a[~valid] == -1).all())
    assert bool((a[valid & clear] == ref[valid & clear]).all())
    assert bool((a[valid] != K - 1).all())
    onehot = torch.zeros_like(T)
    onehot[valid, a[valid]] = 1.0
    assert torch.equal(T, onehot)                      # exact rows, pad columns zero, empty rows all zero
    inertia = ((Z[:, :d].double() - cent.double()[a.clamp_min(0)]) ** 2).sum(1)[valid].sum()
    assert int(stats[2]) == int(valid.sum())
    assert abs(float(stats[0]) - float(inertia)) <= 1e-4 * float(inertia) + 1e-3      # exact fp32 distance to the winner
