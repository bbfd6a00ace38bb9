"""Each CUDA kernel against the plain-PyTorch implementation of the same op (fp32/fp64 reference)."""
import pytest
import torch

from omldm_b200.ops import functional as F
from omldm_b200.oracle.preprocessors import poly_factor_table
from omldm_b200.stream import ring as R

pytestmark = pytest.mark.gpu
DEV = "cuda"


def both(fn):
    """Run fn() once with the torch implementation and once native; fn returns tensors to compare."""
    with F.use_impl("torch"):
        ref = fn()
    nat = fn()
    torch.cuda.synchronize()
    return ref, nat


def staged(kind, d, B, seed=0, n_aux=0):
    return R.generate_batch(kind, d, B, seed=seed, n_aux=n_aux).to(DEV)


def test_native_library_loaded():
    from omldm_b200.ops import _lib
    assert _lib.kernels().omldm_abi_version() == 1


@pytest.mark.parametrize("B,count0,cap", [(100, 0, 16), (4096, 0, 256), (1000, 12345, 256), (37, 8, 4), (20000, 777, 64),
                                          (65536, 65536 * 3, 256)])
def test_holdout_swap(B, count0, cap):
    d = 12
    Z0 = staged(R.KIND_BINARY, d, B, seed=3)
    ring0 = torch.randn(cap, Z0.shape[1], device=DEV)

    def run():
        Z, ring = Z0.clone(), ring0.clone()
        F.holdout_swap(Z, ring, d, count0)
        return Z, ring
    (Zr, rr), (Zn, rn) = both(run)
    assert torch.equal(Zr, Zn) and torch.equal(rr, rn)


@pytest.mark.parametrize("kind,variant", [(F.KIND_PA, 0), (F.KIND_PA, 1), (F.KIND_PA, 2), (F.KIND_PA_REG, 1),
                                          (F.KIND_SVM, 0)])
@pytest.mark.parametrize("d", [28, 434])
def test_linear_exact(kind, variant, d):
    B = 300
    Z = staged(R.KIND_REGRESSION if kind == F.KIND_PA_REG else R.KIND_BINARY, d, B, seed=5)
    Z[7, d] = 0.0   # an empty row must be skipped

    def run():
        w = torch.zeros(d + 1, device=DEV)
        st = torch.zeros(4, dtype=torch.float64, device=DEV)
        F.linear_exact(Z, d, w, st, kind, variant, 0.5, 0.05, 0.01)
        return w, st
    (wr, sr), (wn, sn) = both(run)
    assert torch.allclose(wr, wn, rtol=2e-3, atol=2e-4)
    assert torch.allclose(sr, sn, rtol=2e-3, atol=1e-3)
    assert sn[2].item() == B - 1


@pytest.mark.parametrize("kind", [F.KIND_PA, F.KIND_PA_REG, F.KIND_SVM])
@pytest.mark.parametrize("d,B", [(28, 1000), (434, 4096)])
def test_linear_minibatch(kind, d, B):
    Z = staged(R.KIND_REGRESSION if kind == F.KIND_PA_REG else R.KIND_BINARY, d, B, seed=6)
    w0 = 0.01 * torch.randn(d + 1, device=DEV)

    def run():
        w = w0.clone()
        st = torch.zeros(4, dtype=torch.float64, device=DEV)
        sc = F.linear_minibatch_scratch(d, DEV)
        for _ in range(2):
            F.linear_minibatch(Z, d, w, st, sc, kind, 1, 0.5, 0.05, 0.01)
        return w, st, sc
    (wr, sr, _), (wn, sn, scn) = both(run)
    assert torch.allclose(wr, wn, rtol=1e-3, atol=1e-5)
    assert torch.allclose(sr, sn, rtol=1e-4)
    assert float(scn.abs().sum()) == 0.0     # the kernel re-zeroes its scratch


def test_linear_predict():
    d, B = 100, 513
    Z = staged(R.KIND_BINARY, d, B)
    w = torch.randn(d + 1, device=DEV)
    ref, nat = both(lambda: F.linear_predict(Z, d, w))
    assert torch.allclose(ref, nat, rtol=1e-4, atol=1e-4)


def test_stage_rows_and_densify():
    d, B = 50, 64
    raw = torch.randn(B, d + 1, device=DEV)
    ref, nat = both(lambda: F.stage_rows(raw, d))
    assert torch.equal(ref, nat)
    nnz = 7
    indptr = torch.arange(0, (B + 1) * nnz, nnz, dtype=torch.int32, device=DEV)
    indices = torch.randint(0, d, (B * nnz,), dtype=torch.int32, device=DEV)
    values = torch.randn(B * nnz, device=DEV)
    tg = torch.randn(B, device=DEV)
    (Zr, sr), (Zn, sn) = both(lambda: F.densify(indptr, indices, values, tg, d))
    assert torch.allclose(Zr, Zn, atol=1e-5) and torch.allclose(sr, sn, rtol=1e-4)


@pytest.mark.parametrize("degree", [2, 3])
def test_poly_features(degree):
    d, B = 9, 200
    Z = staged(R.KIND_REGRESSION, d, B)
    Z[3, d] = 0.0
    tab = torch.from_numpy(poly_factor_table(d, degree)).to(DEV)
    ref, nat = both(lambda: F.poly_features(Z, d, tab, tab.shape[0]))
    assert torch.allclose(ref, nat, rtol=1e-5, atol=1e-6)


def test_standard_scaler_and_minmax():
    d, B = 45, 3000
    batches = [3.0 * staged(R.KIND_REGRESSION, d, B, seed=s) + 1.5 for s in range(3)]
    for b in batches:
        b[:, d] = 1.0

    def run_std():
        st = torch.zeros(1 + 2 * d, dtype=torch.float64, device=DEV)
        sc = torch.zeros(3 + 4 * d, dtype=torch.float64, device=DEV)
        outs = []
        for b in batches:
            z = b.clone()
            F.standard_scaler(z, d, st, sc, True)
            outs.append(z)
        return torch.cat(outs), st
    (zr, sr), (zn, sn) = both(run_std)
    assert torch.allclose(sr, sn, rtol=1e-6, atol=1e-6)
    assert torch.allclose(zr, zn, rtol=1e-3, atol=1e-4)

    def run_mm():
        st = torch.cat([torch.full((d,), float("inf")), torch.full((d,), float("-inf"))]).to(DEV)
        z = batches[0].clone()
        F.minmax_scaler(z, d, st, True)
        return z, st
    (zr, sr), (zn, sn) = both(run_mm)
    assert torch.equal(sr, sn) and torch.allclose(zr, zn, rtol=1e-5, atol=1e-6)


def test_kmeans_kernels():
    d, K, B = 128, 256, 2048
    Z = staged(R.KIND_CLUSTERS, d, B, seed=9, n_aux=K)
    ld = Z.shape[1]

    def run():
        sums = torch.zeros(K, ld, device=DEV)
        st = torch.zeros(4, dtype=torch.float64, device=DEV)
        F.kmeans_exact(Z[:600], d, sums, K, st)
        cent = torch.zeros(K, ld, device=DEV)
        cn = torch.zeros(K, device=DEV)
        F.kmeans_finalize(sums, None, K, d, cent, cn)
        S = (Z[:, :d].double() @ cent[:, :d].double().t()).float()
        T = torch.empty(B, F.pad32(K), device=DEV)
        asg = torch.empty(B, dtype=torch.int32, device=DEV)
        F.kmeans_assign(Z, d, S, cn, K, T, st, asg)
        return sums, cent, cn, T, st
    ref, nat = both(run)
    for r, n in zip(ref, nat):
        assert torch.allclose(r, n, rtol=2e-3, atol=2e-3)


def test_mcpa_coef_and_helpers():
    d, C, B = 300, 10, 777
    Z = staged(R.KIND_MULTICLASS, d, B, seed=2, n_aux=C)
    W = 0.1 * torch.randn(C, Z.shape[1], device=DEV)
    S = (Z[:, :d + 1] @ W[:, :d + 1].t()).contiguous()
    S4 = torch.zeros(B, 12, device=DEV)
    S4[:, :C] = S

    def run():
        sq = F.row_sqnorm(Z, d + 1)
        T = torch.empty(B, 32, device=DEV)
        st = torch.zeros(4, dtype=torch.float64, device=DEV)
        F.mcpa_coef(Z, d, S4, C, sq, 1, 0.5, T, st)
        act = st[3:4].float()
        F.scale_rows(T, C, act, reciprocal=True)
        return sq, T, st
    ref, nat = both(run)
    for r, n in zip(ref, nat):
        assert torch.allclose(r, n, rtol=1e-4, atol=1e-6)


def test_sq_residual_mirror_axpy():
    d, B = 1000, 1500
    Z = staged(R.KIND_REGRESSION, d, B)
    wt = torch.randn(d + 2, device=DEV) * 0.03

    def run():
        st = torch.zeros(4, dtype=torch.float64, device=DEV)
        F.sq_residual(Z, d, wt, st)
        A = torch.arange(100 * 128, device=DEV, dtype=torch.float32).reshape(100, 128).clone()
        F.mirror_upper(A, 100)
        y = torch.ones(1000, device=DEV)
        F.axpy(y, torch.arange(1000, device=DEV, dtype=torch.float32), 0.5)
        return st, A, y
    ref, nat = both(run)
    assert torch.allclose(ref[0], nat[0], rtol=1e-4)
    assert torch.equal(ref[1], nat[1]) and torch.equal(ref[2], nat[2])


def test_hoeffding_tree_kernels():
    d, C, B = 20, 3, 3000
    Z = staged(R.KIND_MULTICLASS, d, B, seed=4, n_aux=C)
    Z[11, d] = 0.0
    cap = 7
    feature = torch.tensor([3, 7, -1, -1, -1, -1, -1], dtype=torch.int32, device=DEV)
    threshold = torch.tensor([0.1, -0.2, 0, 0, 0, 0, 0], device=DEV)
    left = torch.tensor([1, 3, -1, -1, -1, -1, -1], dtype=torch.int32, device=DEV)
    right = torch.tensor([2, 4, -1, -1, -1, -1, -1], dtype=torch.int32, device=DEV)
    slot = torch.arange(cap, dtype=torch.int32, device=DEV)

    def run():
        cnt = torch.zeros(cap, C, device=DEV)
        cnt[2, 1] = 5.0   # a prior, so the prequential prediction is defined for that leaf
        sm = torch.zeros(cap, C, d, device=DEV)
        sq = torch.zeros(cap, C, d, device=DEV)
        st = torch.zeros(4, dtype=torch.float64, device=DEV)
        leaf = torch.empty(B, dtype=torch.int32, device=DEV)
        F.ht_accumulate(Z, d, feature, threshold, left, right, slot, C, cnt, sm, sq, st, leaf)
        routed = F.ht_route(Z, feature, threshold, left, right)
        return cnt, sm, sq, st[2:3], leaf, routed
    ref, nat = both(run)
    assert torch.equal(ref[0], nat[0]) and torch.equal(ref[4], nat[4]) and torch.equal(ref[5], nat[5])
    assert torch.allclose(ref[1], nat[1], rtol=1e-4, atol=1e-3) and torch.allclose(ref[2], nat[2], rtol=1e-4, atol=1e-3)
    assert ref[3].item() == nat[3].item() == B - 1
    assert set(nat[5].unique().tolist()) == {2, 3, 4}


@pytest.mark.parametrize("kind", [F.KIND_PA_REG, F.KIND_PA])
@pytest.mark.parametrize("d_in,degree", [(28, 2), (6, 3)])
def test_fused_poly_scaler_linear_equals_separate_kernels(kind, d_in, degree):
    B = 4000
    tab = torch.from_numpy(poly_factor_table(d_in, degree)).to(DEV)
    d_out = tab.shape[0]
    batches = [staged(R.KIND_REGRESSION if kind == F.KIND_PA_REG else R.KIND_BINARY, d_in, B, seed=s) for s in range(3)]
    batches[1][5, d_in] = 0.0

    def run(fused):
        st = torch.zeros(1 + 2 * d_out, dtype=torch.float64, device=DEV)
        sc = torch.zeros(3 + 4 * d_out, dtype=torch.float64, device=DEV)
        w = torch.zeros(d_out + 1, device=DEV)
        stats = torch.zeros(4, dtype=torch.float64, device=DEV)
        ls = F.linear_minibatch_scratch(d_out, DEV)
        for Z in batches:
            if fused:
                F.fused_poly_scaler_linear(Z, d_in, tab, d_out, st, sc, w, stats, ls, kind, 1, 0.5, 0.05, 0.01)
            else:
                Zp = F.poly_features(Z, d_in, tab, d_out)
                F.standard_scaler(Zp, d_out, st, sc, True)
                F.linear_minibatch(Zp, d_out, w, stats, ls, kind, 1, 0.5, 0.05, 0.01)
        torch.cuda.synchronize()
        return w, stats, st
    w1, s1, st1 = run(True)
    w0, s0, st0 = run(False)
    with F.use_impl("torch"):
        wt, stt, _ = run(False)
    assert torch.allclose(st1, st0, rtol=1e-9, atol=1e-9)
    assert torch.allclose(w1, w0, rtol=2e-3, atol=2e-5) and torch.allclose(w1, wt, rtol=5e-3, atol=5e-5)
    assert torch.allclose(s1, s0, rtol=1e-3)
