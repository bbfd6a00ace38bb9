"""GEMM whose epilogue is the reduce-scatter (umma_gemm_scatter_kernel) on ONE GPU: the "peers" are separate local
buffers, which checks the per-strip choice of the output tensor map and the TMA reduce-add into it without needing a
second device.  First hardware run pending (written after round 1's GPU budget was spent)."""
import pytest
import torch

from omldm_b200.ops import functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_peers,d,K,ks", [(4, 510, 4096, 4), (8, 1022, 2048, 7), (2, 254, 1024, 1)])
def test_rows_land_in_their_owners_buffers(n_peers, d, K, ks):
    ld = F.pad32(d + 2)
    rows = ld // n_peers
    assert rows % 32 == 0
    Z = torch.randn(K, ld, device="cuda")
    Z[:, d + 2:] = 0.0
    n = d + 2
    bufs = [torch.zeros(ld, ld, device="cuda") for _ in range(n_peers)]
    for _ in range(2):                                       # accumulates: two calls = twice the product
        F.gemm_outer_scatter(Z, Z, [b.data_ptr() for b in bufs], rows, ld, ksplit=ks, upper_only=True, m=n, n=n)
    torch.cuda.synchronize()
    ref = 2.0 * (Z[:, :n].double().t() @ Z[:, :n].double()).float()
    bound = 2.0 * 2.0 ** -9 * (Z[:, :n].abs().double().t() @ Z[:, :n].abs().double()).float()   # tf32 products
    for p, b in enumerate(bufs):
        lo, hi = p * rows, min((p + 1) * rows, n)
        outside = b.clone()
        outside[lo:hi] = 0.0
        assert float(outside.abs().max()) == 0.0, f"peer {p} received rows it does not own"
        if hi <= lo:
            continue
        got = torch.triu(b[:n, :n])[lo:hi]                    # only tiles on / above the diagonal are computed
        want = torch.triu(ref)[lo:hi]
        tile_ok = torch.triu(torch.ones(n, n, device="cuda"))[lo:hi] > 0
        err = ((got - want).abs() * tile_ok)
        assert bool((err <= bound[lo:hi] + 1e-3).all()), f"peer {p}: max err {float(err.max())}"
