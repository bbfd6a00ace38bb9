"""The REAL kernel sources (csrc/*.cu without inline PTX) executed on the CPU through tools/cuda_emu - one host
thread per CUDA thread, real barriers / shuffles / atomics - and compared with the plain-PyTorch implementation
of the same op.  This is the CPU-only counterpart of tests/test_kernels_gpu.py (smaller shapes: every block is a
few hundred host threads) and, with OMLDM_EMU_ASAN=1, the memory-safety check of the kernel code
(`tests/test_emu_asan.py` runs this file under AddressSanitizer in a subprocess)."""
import os

import pytest
import torch

from omldm_b200.ops import _lib
from omldm_b200.ops import functional as F
from omldm_b200.oracle.preprocessors import poly_factor_table
from omldm_b200.stream import ring as R

DEV = "cpu"


@pytest.fixture(scope="module", autouse=True)
def emulator():
    if torch.cuda.is_available():
        pytest.skip("the emulator is for machines without a GPU")
    _lib.use_emulator(asan=bool(os.environ.get("OMLDM_EMU_ASAN")))
    yield
    _lib.stop_emulator()


def both(fn):
    with F.use_impl("torch"):
        ref = fn()
    with F.use_impl("emu"):
        nat = fn()
    return ref, nat


def staged(kind, d, B, seed=0, n_aux=0):
    return R.generate_batch(kind, d, B, seed=seed, n_aux=n_aux)


def test_emulated_library_is_loaded():
    assert _lib.kernels().omldm_is_emulator() == 1 and _lib.kernels().omldm_abi_version() == 1


@pytest.mark.parametrize("B,count0,cap", [(100, 0, 16), (300, 12345, 16), (37, 8, 4), (700, 777, 8)])
def test_holdout_swap(B, count0, cap):
    d = 12
    Z0 = staged(R.KIND_BINARY, d, B, seed=3)
    ring0 = torch.randn(cap, Z0.shape[1])

    def run():
        Z, ring = Z0.clone(), ring0.clone()
        F.holdout_swap(Z, ring, d, count0)
        return Z, ring
    (Zr, rr), (Zn, rn) = both(run)
    assert torch.equal(Zr, Zn) and torch.equal(rr, rn)


@pytest.mark.parametrize("B,count0,cap,ld", [(100, 0, 16, 32), (300, 12345, 16, 32), (37, 8, 4, 64), (700, 777, 8, 512),
                                             (2000, 4321, 16, 1024), (50, 3, 256, 32)])
def test_holdout_exchange_keeps_the_trained_set_and_the_ring(B, count0, cap, ld):
    """The exchange form of the holdout (one swap per ring slot, for learners that only sum over the batch) against the
    ordered shift: the ring ends up identical, the batch holds the same MULTISET of rows, the device-resident record
    count advances by B - for narrow rows (several slots per CTA), 2 KB rows (8-byte lanes) and wide rows."""
    g = torch.Generator().manual_seed(B + cap)
    Z0 = torch.randn(B, ld, generator=g)
    ring0 = torch.randn(cap, ld, generator=g)
    out = []
    with F.use_impl("emu"):
        for ordered in (True, False):
            Z, ring = Z0.clone(), ring0.clone()
            cnt = torch.tensor([count0, 0], dtype=torch.int64)
            F.holdout_swap_dev(Z, ring, ld - 2, cnt, ordered=ordered)
            assert int(cnt[0]) == count0 + B and int(cnt[1]) == 0
            out.append((Z, ring))
    (Zo, ro), (Zx, rx) = out
    Zref, rref = Z0.clone(), ring0.clone()
    F.holdout_swap(Zref, rref, ld - 2, count0)                      # PyTorch restatement of the sequential rule
    assert torch.equal(Zo, Zref) and torch.equal(ro, rref)
    assert torch.equal(rx, ro)
    key = lambda t: sorted(map(tuple, t.tolist()))
    assert key(Zx) == key(Zo)


@pytest.mark.parametrize("kind,variant", [(F.KIND_PA, 1), (F.KIND_PA_REG, 1), (F.KIND_SVM, 0)])
@pytest.mark.parametrize("d", [28, 70])
def test_linear_exact(kind, variant, d):
    B = 60
    Z = staged(R.KIND_REGRESSION if kind == F.KIND_PA_REG else R.KIND_BINARY, d, B, seed=5)
    Z[7, d] = 0.0

    def run():
        w = torch.zeros(d + 1)
        st = torch.zeros(4, dtype=torch.float64)
        F.linear_exact(Z, d, w, st, kind, variant, 0.5, 0.05, 0.01)
        return w, st
    (wr, sr), (wn, sn) = both(run)
    assert torch.allclose(wr, wn, rtol=2e-3, atol=2e-4) and torch.allclose(sr, sn, rtol=2e-3, atol=1e-3)
    assert sn[2].item() == B - 1


@pytest.mark.parametrize("kind", [F.KIND_PA, F.KIND_PA_REG, F.KIND_SVM])
def test_linear_minibatch_and_static_variant(kind):
    d, B = 45, 300
    Z = staged(R.KIND_REGRESSION if kind == F.KIND_PA_REG else R.KIND_BINARY, d, B, seed=6)
    w0 = 0.01 * torch.randn(d + 1)

    def run():
        w = w0.clone()
        st = torch.zeros(4, dtype=torch.float64)
        sc = F.linear_minibatch_scratch(d, DEV)
        for _ in range(2):
            F.linear_minibatch(Z, d, w, st, sc, kind, 1, 0.5, 0.05, 0.01)
        return w, st, sc
    (wr, sr, _), (wn, sn, scn) = both(run)
    assert torch.allclose(wr, wn, rtol=1e-3, atol=1e-5) and torch.allclose(sr, sn, rtol=1e-4)
    assert float(scn.abs().sum()) == 0.0
    # the no-allocation entry point (fixed scratch inside the library) gives the same update, twice in a row
    import ctypes
    w2 = w0.clone()
    st2 = torch.zeros(4, dtype=torch.float64)
    desc = F.LinearDesc(w2.data_ptr(), st2.data_ptr(), d, kind, 1, 0.5, 0.05, 0.01)
    k = _lib.kernels()
    for _ in range(2):
        rc = k.omldm_linear_minibatch_static(_lib.ptr(Z), ctypes.c_int(Z.stride(0)), ctypes.c_int(B), ctypes.byref(desc),
                                             ctypes.c_int(1), _lib.stream_ptr())
        assert rc == 0, k.omldm_last_error()
    assert torch.allclose(w2, wn, rtol=1e-4, atol=1e-5) and torch.allclose(st2, sn, rtol=1e-6)
    big = F.LinearDesc(w2.data_ptr(), st2.data_ptr(), 4096, kind, 1, 0.5, 0.05, 0.01)     # beyond the documented limit
    assert k.omldm_linear_minibatch_static(_lib.ptr(Z), ctypes.c_int(Z.stride(0)), ctypes.c_int(B), ctypes.byref(big),
                                           ctypes.c_int(1), _lib.stream_ptr()) != 0


def test_linear_predict_stage_rows_densify():
    d, B = 50, 64
    Z = staged(R.KIND_BINARY, d, B)
    w = torch.randn(d + 1)
    ref, nat = both(lambda: F.linear_predict(Z, d, w))
    assert torch.allclose(ref, nat, rtol=1e-4, atol=1e-4)
    raw = torch.randn(B, d + 1)
    ref, nat = both(lambda: F.stage_rows(raw, d))
    assert torch.equal(ref, nat)
    nnz = 7
    indptr = torch.arange(0, (B + 1) * nnz, nnz, dtype=torch.int32)
    indices = torch.randint(0, d, (B * nnz,), dtype=torch.int32)
    indices[5] = d + 3                       # out of range: must be dropped, not written
    values = torch.randn(B * nnz)
    tg = torch.randn(B)
    (Zr, sr), (Zn, sn) = both(lambda: F.densify(indptr, indices, values, tg, d))
    assert torch.allclose(Zr, Zn, atol=1e-5) and torch.allclose(sr, sn, rtol=1e-4)


@pytest.mark.parametrize("d", [1, 5, 28, 63])
def test_stage_rows_f16(d):
    B = 33
    w = F.pack_wire_f16(torch.randn(B, d + 1), d)
    ref, nat = both(lambda: F.stage_rows_f16(w, d))
    assert torch.equal(ref, nat)
    out = torch.full((B, F.pad32(d + 2)), float("nan"))
    with F.use_impl("emu"):
        F.stage_rows_f16(w, d, out=out)
    assert not bool(torch.isnan(out).any())


@pytest.mark.parametrize("degree", [2, 3])
def test_poly_features(degree):
    d, B = 6, 80
    Z = staged(R.KIND_REGRESSION, d, B)
    Z[3, d] = 0.0
    tab = torch.from_numpy(poly_factor_table(d, degree))
    ref, nat = both(lambda: F.poly_features(Z, d, tab, tab.shape[0]))
    assert torch.allclose(ref, nat, rtol=1e-5, atol=1e-6)


def test_standard_scaler_and_minmax():
    d, B = 21, 400
    batches = [3.0 * staged(R.KIND_REGRESSION, d, B, seed=s) + 1.5 for s in range(3)]
    for b in batches:
        b[:, d] = 1.0

    def run_std():
        st = torch.zeros(1 + 2 * d, dtype=torch.float64)
        sc = torch.zeros(3 + 4 * d, dtype=torch.float64)
        outs = []
        for b in batches:
            z = b.clone()
            F.standard_scaler(z, d, st, sc, True)
            outs.append(z)
        return torch.cat(outs), st
    (zr, sr), (zn, sn) = both(run_std)
    assert torch.allclose(sr, sn, rtol=1e-6, atol=1e-6) and torch.allclose(zr, zn, rtol=1e-3, atol=1e-4)

    def run_mm():
        st = torch.cat([torch.full((d,), float("inf")), torch.full((d,), float("-inf"))])
        z = batches[0].clone()
        F.minmax_scaler(z, d, st, True)
        return z, st
    (zr, sr), (zn, sn) = both(run_mm)
    assert torch.equal(sr, sn) and torch.allclose(zr, zn, rtol=1e-5, atol=1e-6)


def test_kmeans_kernels():
    d, K, B = 16, 8, 200
    Z = staged(R.KIND_CLUSTERS, d, B, seed=9, n_aux=K)
    ld = Z.shape[1]

    def run():
        sums = torch.zeros(K, ld)
        st = torch.zeros(4, dtype=torch.float64)
        F.kmeans_exact(Z[:60], d, sums, K, st)
        cent = torch.zeros(K, ld)
        cn = torch.zeros(K)
        F.kmeans_finalize(sums, None, K, d, cent, cn)
        S = (Z[:, :d].double() @ cent[:, :d].double().t()).float()
        T = torch.empty(B, F.pad32(K))
        asg = torch.empty(B, dtype=torch.int32)
        F.kmeans_assign(Z, d, S, cn, K, T, st, asg)
        return sums, cent, cn, T, st
    ref, nat = both(run)
    for r, n in zip(ref, nat):
        assert torch.allclose(r, n, rtol=2e-3, atol=2e-3)


def test_mcpa_coef_and_helpers():
    d, C, B = 40, 10, 77
    Z = staged(R.KIND_MULTICLASS, d, B, seed=2, n_aux=C)
    W = 0.1 * torch.randn(C, Z.shape[1])
    S4 = torch.zeros(B, 12)
    S4[:, :C] = Z[:, :d + 1] @ W[:, :d + 1].t()

    def run():
        sq = F.row_sqnorm(Z, d + 1)
        T = torch.empty(B, 32)
        st = torch.zeros(4, dtype=torch.float64)
        F.mcpa_coef(Z, d, S4, C, sq, 1, 0.5, T, st)
        act = st[3:4].float()
        F.scale_rows(T, C, act, reciprocal=True)
        return sq, T, st
    ref, nat = both(run)
    for r, n in zip(ref, nat):
        assert torch.allclose(r, n, rtol=1e-4, atol=1e-6)


def test_sq_residual_mirror_axpy():
    d, B = 100, 150
    Z = staged(R.KIND_REGRESSION, d, B)
    wt = torch.randn(d + 2) * 0.03

    def run():
        st = torch.zeros(4, dtype=torch.float64)
        F.sq_residual(Z, d, wt, st)
        A = torch.arange(40 * 64, dtype=torch.float32).reshape(40, 64).clone()
        F.mirror_upper(A, 40)
        y = torch.ones(300)
        F.axpy(y, torch.arange(300, dtype=torch.float32), 0.5)
        return st, A, y
    ref, nat = both(run)
    assert torch.allclose(ref[0], nat[0], rtol=1e-4)
    assert torch.equal(ref[1], nat[1]) and torch.equal(ref[2], nat[2])


def test_hoeffding_tree_kernels():
    d, C, B = 10, 3, 300
    Z = staged(R.KIND_MULTICLASS, d, B, seed=4, n_aux=C)
    Z[11, d] = 0.0
    cap = 7
    feature = torch.tensor([3, 7, -1, -1, -1, -1, -1], dtype=torch.int32)
    threshold = torch.tensor([0.1, -0.2, 0, 0, 0, 0, 0])
    left = torch.tensor([1, 3, -1, -1, -1, -1, -1], dtype=torch.int32)
    right = torch.tensor([2, 4, -1, -1, -1, -1, -1], dtype=torch.int32)
    slot = torch.arange(cap, dtype=torch.int32)

    def run():
        cnt = torch.zeros(cap, C)
        cnt[2, 1] = 5.0
        sm = torch.zeros(cap, C, d)
        sq = torch.zeros(cap, C, d)
        st = torch.zeros(4, dtype=torch.float64)
        leaf = torch.empty(B, dtype=torch.int32)
        F.ht_accumulate(Z, d, feature, threshold, left, right, slot, C, cnt, sm, sq, st, leaf)
        routed = F.ht_route(Z, feature, threshold, left, right)
        return cnt, sm, sq, st[2:3], leaf, routed
    ref, nat = both(run)
    assert torch.equal(ref[0], nat[0]) and torch.equal(ref[4], nat[4]) and torch.equal(ref[5], nat[5])
    assert torch.allclose(ref[1], nat[1], rtol=1e-4, atol=1e-3) and torch.allclose(ref[2], nat[2], rtol=1e-4, atol=1e-3)
    assert ref[3].item() == nat[3].item() == B - 1


@pytest.mark.parametrize("kind", [F.KIND_PA_REG, F.KIND_PA])
@pytest.mark.parametrize("d_in,degree", [(8, 2), (4, 3)])
def test_fused_poly_scaler_linear_equals_separate_kernels(kind, d_in, degree):
    B = 300
    tab = torch.from_numpy(poly_factor_table(d_in, degree))
    d_out = tab.shape[0]
    batches = [staged(R.KIND_REGRESSION if kind == F.KIND_PA_REG else R.KIND_BINARY, d_in, B, seed=s) for s in range(2)]
    batches[1][5, d_in] = 0.0

    def run(fused):
        st = torch.zeros(1 + 2 * d_out, dtype=torch.float64)
        sc = torch.zeros(3 + 4 * d_out, dtype=torch.float64)
        w = torch.zeros(d_out + 1)
        stats = torch.zeros(4, dtype=torch.float64)
        ls = F.linear_minibatch_scratch(d_out, DEV)
        for Z in batches:
            if fused:
                F.fused_poly_scaler_linear(Z, d_in, tab, d_out, st, sc, w, stats, ls, kind, 1, 0.5, 0.05, 0.01)
            else:
                Zp = F.poly_features(Z, d_in, tab, d_out)
                F.standard_scaler(Zp, d_out, st, sc, True)
                F.linear_minibatch(Zp, d_out, w, stats, ls, kind, 1, 0.5, 0.05, 0.01)
        return w, stats, st
    with F.use_impl("emu"):
        w1, s1, st1 = run(True)
        w0, s0, st0 = run(False)
    with F.use_impl("torch"):
        wt, stt, _ = run(False)
    assert torch.allclose(st1, st0, rtol=1e-9, atol=1e-9)
    assert torch.allclose(w1, w0, rtol=2e-3, atol=2e-5) and torch.allclose(w1, wt, rtol=5e-3, atol=5e-5)
    assert torch.allclose(s1, s0, rtol=1e-3)


# ---- kernels written after the round's GPU budget was spent: this is their first execution of any kind ----------
def _sparse_slot(i, d, B, nnz, C, unique):
    g = torch.Generator().manual_seed(100 + i)
    if unique:
        idx = torch.stack([torch.randperm(d, generator=g)[:nnz] for _ in range(B)]).int()
    else:
        idx = torch.randint(0, d, (B, nnz), generator=g).int()
    val = torch.randn(B, nnz, generator=g)
    y = torch.randint(0, C, (B,), generator=g).float()
    return idx.reshape(-1), val.reshape(-1), y


@pytest.mark.parametrize("unique", [True, False])
@pytest.mark.parametrize("mean", [True, False])
@pytest.mark.parametrize("class_minor", [False, True])
def test_sparse_mcpa_step(unique, mean, class_minor):
    d, B, nnz, C = 500, 50, 12, 6
    wt = torch.empty((d + 1) * 16) if class_minor else None     # scores gathered from the class-minor copy of W
    ldw = F.pad32(d + 2)
    W0 = 0.05 * torch.randn(C, ldw)
    W0[:, d + 1:] = 0.0
    batches = [_sparse_slot(i, d, B, nnz, C, unique) for i in range(3)]

    def run_emu():
        W = W0.clone()
        st = torch.zeros(4, dtype=torch.float64)
        coef = torch.empty(B, 4)
        for idx, val, y in batches:
            rows = F.sparse_rows(idx, val, y, B, nnz)
            rows[4] = 0.0                                    # an empty row is skipped
            prev = st[3].clone()
            F.mcpa_sparse_step(rows, nnz, W, d, C, 1, 0.5, coef, st, prev, mean, wt)
        return W, st

    def run_ref():
        """The dense MultiClassPA mini-batch step on the densified rows, with the sparse path's norm (sum v^2 + 1)."""
        W = W0.clone()
        st = torch.zeros(4, dtype=torch.float64)
        for idx, val, y in batches:
            rows = F.sparse_rows(idx, val, y, B, nnz)
            rows[4] = 0.0
            Z = F.sparse_rows_to_dense(rows, nnz, d)
            col = rows[:, :nnz].contiguous().view(torch.int32)
            v = rows[:, nnz:2 * nnz]
            sq = torch.where((col >= 0) & (col < d), v * v, torch.zeros_like(v)).sum(1) + 1.0
            S = torch.zeros(B, 8)
            S[:, :C] = Z[:, :d + 1] @ W[:, :d + 1].t()
            T = torch.empty(B, 32)
            prev = st[3].clone()
            F.mcpa_coef(Z, d, S, C, sq, 1, 0.5, T, st)
            if mean:
                F.scale_rows(T, C, (st[3] - prev).float().reshape(1), reciprocal=True)
            W[:, :d + 1] += T[:, :C].t() @ Z[:, :d + 1]
        return W, st
    with F.use_impl("emu"):
        We, se = run_emu()
    with F.use_impl("torch"):
        Wr, sr = run_ref()
    assert torch.allclose(We, Wr, rtol=1e-4, atol=1e-6)
    assert torch.allclose(se, sr, rtol=1e-5) and se[2].item() == 3 * (B - 1)


@pytest.mark.parametrize("blocks", [1, 3])
@pytest.mark.parametrize("n,m", [(5, 40), (33, 300), (130, 900)])
def test_ridge_cg_kernel(n, m, blocks):
    g = torch.Generator().manual_seed(n)
    X = torch.randn(m, n, generator=g) * (1.0 + 3.0 * torch.rand(n, generator=g))
    y = X @ torch.randn(n, generator=g) + 0.1 * torch.randn(m, generator=g)
    ld = F.pad32(n + 1)
    W = torch.zeros(ld, ld)
    W[:n, :n] = X.t() @ X
    W[:n, n] = X.t() @ y
    ref = torch.linalg.solve(W[:n, :n].double() + torch.eye(n, dtype=torch.float64), W[:n, n].double()).float()
    with F.use_impl("emu"):
        x = torch.zeros(n)
        work, scal, info = F.ridge_cg_workspace(n, DEV, 64)
        F.ridge_cg(W, n, 1.0, x, work, scal, info, max_iter=64, tol=1e-6, blocks=blocks)
        it1, rel1 = info.tolist()
        assert rel1 <= 1e-6 and 0 < it1 <= 64
        assert torch.allclose(x, ref, rtol=2e-3, atol=2e-4)
        F.ridge_cg(W, n, 1.0, x, work, scal, info, max_iter=64, tol=1e-4, blocks=blocks)     # warm start: already converged
        assert int(info[0]) == 0
    with F.use_impl("torch"):                                                  # same iteration count as the model
        x2 = torch.zeros(n)
        F.ridge_cg(W, n, 1.0, x2, *F.ridge_cg_workspace(n, DEV, 64), max_iter=64, tol=1e-6)
    assert torch.allclose(x, x2, rtol=1e-3, atol=1e-4)


@pytest.mark.parametrize("blocks", [1, 3])
@pytest.mark.parametrize("n,m", [(1, 10), (5, 40), (32, 200), (33, 300), (97, 700), (130, 900)])
def test_ridge_chol_kernel(n, m, blocks):
    """Blocked Cholesky with the right-hand side carried as an extra row (csrc/ridge_chol.cu) against an fp64 solve."""
    g = torch.Generator().manual_seed(n)
    X = torch.randn(m, n, generator=g) * (1.0 + 3.0 * torch.rand(n, generator=g))
    y = X @ torch.randn(n, generator=g) + 0.1 * torch.randn(m, generator=g)
    ld = F.pad32(n + 2)
    W = torch.zeros(ld, ld)
    W[:n, :n] = X.t() @ X
    W[:n, n] = W[n, :n] = X.t() @ y
    W[n, n] = y @ y
    ref = torch.linalg.solve(W[:n, :n].double() + 0.5 * torch.eye(n, dtype=torch.float64), W[:n, n].double()).float()
    with F.use_impl("emu"):
        x = torch.zeros(n)
        info = F.ridge_chol(W.clone(), n, 0.5, x, blocks=blocks)
        assert float(info[0]) == 0.0
        assert torch.allclose(x, ref, rtol=2e-3, atol=2e-4), float((x - ref).abs().max())
        Wbad = W.clone()
        Wbad[n // 2, n // 2] = -5.0                       # not positive definite: reported, never a hang
        info = F.ridge_chol(Wbad, n, 0.0, torch.zeros(n), blocks=blocks)
        assert float(info[0]) == 1.0


def test_f16s_row_kernels():
    """expand_rows_f16s / sq_residual_f16s (csrc/rows_f16.cu) against their PyTorch restatements."""
    for d, B in ((1, 3), (5, 7), (28, 100), (61, 33), (125, 64)):
        raw = torch.randn(B, d + 1)
        Zh = F.pack_f16s(raw, d)
        Zh[::5] = 0                                       # empty rows
        wt = torch.randn(d + 3)
        wt[d + 1:] = -1.0
        ref_Z = F.expand_rows_f16s(Zh, d)
        st_ref = torch.zeros(4, dtype=torch.float64)
        F.sq_residual_f16s(Zh, d, wt, st_ref)
        with F.use_impl("emu"):
            out = torch.full((B, F.pad32(d + 2)), float("nan"))
            F.expand_rows_f16s(Zh, d, out=out)
            assert torch.equal(out, ref_Z), (d, B)
            st = torch.zeros(4, dtype=torch.float64)
            F.sq_residual_f16s(Zh, d, wt, st)
        assert st[2] == st_ref[2] and abs(float(st[0] - st_ref[0])) <= 1e-4 * max(1.0, float(st_ref[0]))
        # the learning-curve sample written by the pass's last block: same ring content as curve_sample after the pass
        with F.use_impl("emu"):
            st2 = torch.zeros(4, dtype=torch.float64)
            ring, cnt = torch.zeros(3, 2, dtype=torch.float64), torch.zeros(2, dtype=torch.int64)
            for k in range(4):                                # wraps around the 3-slot ring
                F.sq_residual_f16s(Zh, d, wt, st2, curve_ring=ring, curve_counter=cnt)
                assert int(cnt[0]) == k + 1 and int(cnt[1]) == 0        # sample taken, completion word reset
                assert ring[k % 3, 0] == st2[0] and ring[k % 3, 1] == st2[2]


# ---- engine-level: the Python glue of the kernel paths that normally only run on a GPU ---------------------------
def _mk_engine(learner, tc=None, pre=None, extra=None, d=12):
    from omldm_b200.engine import Engine
    from omldm_b200.utils.config import JobParameters
    jp = JobParameters(parallelism=1, testSetSize=16)
    jp.extra.update(extra or {})
    e = Engine(jp, device="cpu")
    e.handle_request({"id": 1, "request": "Create", "learner": learner, "preProcessors": pre or [],
                      "trainingConfiguration": dict({"protocol": "Synchronous", "nFeatures": d}, **(tc or {}))})
    return e


def test_engine_sparse_direct_on_emulated_kernels():
    d, B, nnz, C = 800, 40, 10, 5
    slots = []
    for i in range(5):
        idx, val, y = _sparse_slot(i, d, B, nnz, C, unique=False)
        slots.append(torch.cat([idx.view(torch.uint8), val.view(torch.uint8), y.view(torch.uint8)]))
    learner = {"name": "MultiClassPA", "hyperParameters": {"nClasses": C, "C": 0.1}}
    out = []
    for impl in ("torch", "emu"):
        with F.use_impl(impl):
            e = _mk_engine(learner, extra={"sparseDirect": "true"}, d=d)
            for s in slots:
                e.train_sparse(s.clone(), B, nnz, d)
            out.append((e.pipelines[1].learner.W.clone(), e.pipelines[1].learner.counters(), e.holdout.clone()))
    (Wt, ct, ht), (We, ce, he) = out
    assert torch.allclose(Wt, We, rtol=1e-4, atol=1e-6) and torch.equal(ht, he)
    assert ct["fitted"] == ce["fitted"] and abs(ct["loss_sum"] - ce["loss_sum"]) < 1e-4 * ct["loss_sum"]


@pytest.mark.parametrize("learner,kind,pre", [
    ({"name": "PA", "hyperParameters": {"C": 0.1}}, R.KIND_BINARY, None),
    ({"name": "RegressorPA", "hyperParameters": {"C": 0.1, "epsilon": 0.05}}, R.KIND_REGRESSION,
     [{"name": "PolynomialFeatures", "hyperParameters": {"degree": 2}}, {"name": "StandardScaler"}]),
    ({"name": "SVM", "hyperParameters": {"C": 0.1}}, R.KIND_BINARY, [{"name": "MinMaxScaler"}]),
    ({"name": "K-means", "hyperParameters": {"k": 4}}, R.KIND_CLUSTERS, None),
    ({"name": "ORR", "hyperParameters": {"lambda": 1.0, "solver": "cg", "solveEvery": 2}}, R.KIND_REGRESSION, None),
])
def test_engine_pipelines_on_emulated_kernels(learner, kind, pre):
    d, B = 12, 96
    batches = [R.generate_batch(kind, d, B, seed=11, batch_index=i, n_aux=4) for i in range(5)]
    out = []
    for impl in ("torch", "emu"):
        with F.use_impl(impl):
            e = _mk_engine(learner, pre=pre, d=d)
            for b in batches:
                e.train_batch(b.clone(), d)
            l = e.pipelines[1].learner
            if hasattr(l, "solve"):
                l.solve()
            out.append((l.buf.clone() if l.name != "ORR" else l.wt.clone(), l.counters()))
    (bt, ct), (be, ce) = out
    assert ct["fitted"] == ce["fitted"] > 0
    assert torch.allclose(bt, be, rtol=5e-3, atol=5e-4)
    assert abs(ct["loss_sum"] - ce["loss_sum"]) <= 5e-3 * max(1.0, abs(ct["loss_sum"]))


def test_engine_multi_tenant_exact_and_device_count_holdout_on_emulated_kernels():
    """Three exact-mode linear pipelines advance in ONE launch (one warp each), the holdout uses the device-resident
    record count (the graph-replayable variant): same models as three separate torch-path pipelines."""
    d, B = 12, 40
    batches = [R.generate_batch(R.KIND_BINARY, d, B, seed=21, batch_index=i) for i in range(4)]
    specs = [{"name": "PA", "hyperParameters": {"C": 0.1, "trainingMode": "exact"}},
             {"name": "SVM", "hyperParameters": {"C": 0.5, "trainingMode": "exact"}},
             {"name": "PA", "hyperParameters": {"C": 1.0, "updateType": "PA-II", "trainingMode": "exact"}}]
    out = []
    for impl in ("torch", "emu"):
        with F.use_impl(impl):
            e = _mk_engine(specs[0], d=d)
            for j, sp_ in enumerate(specs[1:], start=2):
                e.handle_request({"id": j, "request": "Create", "learner": sp_,
                                  "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
            for b in batches:
                e.train_batch(b.clone(), d)
            out.append(([e.pipelines[i].learner.w.clone() for i in (1, 2, 3)], e.holdout.clone(), e.count,
                        None if e._count_dev is None else int(e._count_dev[0].item())))
    (wt, ht, ct, _), (we, he, ce, cdev) = out
    assert ct == ce == cdev == 4 * B and torch.equal(ht, he)
    for a, b in zip(wt, we):
        assert torch.allclose(a, b, rtol=2e-3, atol=2e-4)


def test_engine_hoeffding_tree_on_emulated_kernels():
    d, B, C = 6, 120, 3
    batches = [R.generate_batch(R.KIND_CLUSTERS, d, B, seed=31, batch_index=i, n_aux=C) for i in range(6)]   # y = cluster id
    probe = R.generate_batch(R.KIND_CLUSTERS, d, 64, seed=31, batch_index=9, n_aux=C)
    out = []
    for impl in ("torch", "emu"):
        with F.use_impl(impl):
            e = _mk_engine({"name": "HT", "hyperParameters": {"nClasses": C, "gracePeriod": 60, "delta": 1e-3, "tieThreshold": 0.2, "maxNodes": 31}}, d=d)
            for b in batches:
                e.train_batch(b.clone(), d)
            l = e.pipelines[1].learner
            l.to_oracle()                                   # pulls the device-grown tree into the host mirror
            out.append((list(l.tree.feature), [round(float(t), 4) for t in l.tree.threshold], l.counters(),
                        l.predict(probe.clone()).clone()))
    (ft, tt, ct, pt), (fe, te, ce, pe) = out
    # the root split is decided by a wide margin; deeper ones can be near-ties between two features, where the
    # summation order of the float atomics may legitimately tip the choice
    assert len(ft) >= 3 and len(fe) >= 3 and ft[0] == fe[0] and abs(tt[0] - te[0]) < 1e-3
    assert ct["fitted"] == ce["fitted"] and float((pt == pe).float().mean()) >= 0.9


def test_ht_split_kernel_matches_the_host_split_test():
    """On-device Hoeffding split test (ht_split_eval / ht_split_apply) against the fp64 host oracle on the same leaf
    statistics: same decision, feature and threshold for the root, children seeded with the same class priors, node
    numbering in ascending leaf order."""
    import numpy as np
    from omldm_b200.oracle import learners as O
    d, C, B = 7, 3, 600
    Z = staged(R.KIND_CLUSTERS, d, B, seed=41, n_aux=C)
    cap = 15
    for grace, expect_split in ((200, True), (5000, False)):
        hyper = {"nClasses": C, "gracePeriod": grace, "delta": 1e-3, "tieThreshold": 0.2}
        feature = torch.full((cap,), -1, dtype=torch.int32)
        threshold = torch.zeros(cap)
        left = torch.full((cap,), -1, dtype=torch.int32)
        right = torch.full((cap,), -1, dtype=torch.int32)
        slot = torch.arange(cap, dtype=torch.int32)
        cnt, sm, sq = torch.zeros(cap, C), torch.zeros(cap, C, d), torch.zeros(cap, C, d)
        st = torch.zeros(4, dtype=torch.float64)
        with F.use_impl("torch"):
            F.ht_accumulate(Z, d, feature, threshold, left, right, slot, C, cnt, sm, sq, st)
        # host oracle on the same statistics
        t = O.HoeffdingTree(hyper)
        t.ensure_init(d)
        lf = O._Leaf(C, d)
        lf.counts = cnt[0].double().numpy().copy()
        n = np.maximum(lf.counts[:, None], 0.0) * np.ones((1, d))
        s_, q_ = sm[0].double().numpy(), sq[0].double().numpy()
        lf.n = n
        lf.mean = np.divide(s_, n, out=np.zeros_like(s_), where=n > 0)
        lf.m2 = np.maximum(q_ - n * lf.mean ** 2, 0.0)
        t.leaves[0] = lf
        if lf.counts.sum() >= grace:
            t._try_split(0, lf)
        # device
        n_nodes = torch.ones(1, dtype=torch.int32)
        checked = torch.zeros(cap)
        decision = torch.zeros(cap, 4, dtype=torch.int32)
        n_splits = torch.zeros(1, dtype=torch.int32)
        cnt_before = cnt.clone()
        with F.use_impl("emu"):
            F.ht_split(feature, threshold, left, right, n_nodes, C, d, cnt, sm, sq, checked, float(grace), 1e-3, 0.2, 10,
                       decision, n_splits)
        assert (int(n_splits) == 1) == expect_split == (len(t.feature) == 3)
        if not expect_split:
            assert int(n_nodes) == 1 and int(feature[0]) == -1 and torch.equal(cnt, cnt_before)
            continue
        assert int(n_nodes) == 3 and int(feature[0]) == t.feature[0]
        assert abs(float(threshold[0]) - t.threshold[0]) < 1e-3 * max(1.0, abs(t.threshold[0]))
        assert (int(left[0]), int(right[0])) == (1, 2) and feature[1:3].tolist() == [-1, -1]
        for ch in (1, 2):
            assert np.allclose(cnt[ch].numpy(), t.leaves[ch].counts, rtol=2e-3, atol=1e-2)
            assert abs(float(checked[ch]) - float(cnt[ch].sum())) < 1e-3
        assert float(checked[0]) == float(cnt_before[0].sum())
        # a second call right away: nothing has changed since the last check, nothing splits
        with F.use_impl("emu"):
            F.ht_split(feature, threshold, left, right, n_nodes, C, d, cnt, sm, sq, checked, float(grace), 1e-3, 0.2, 10,
                       decision, n_splits)
        assert int(n_splits) == 0 and int(n_nodes) == 3


@pytest.mark.parametrize("d,B", [(1, 1), (3, 2), (33, 5), (31, 33), (65, 1)])
def test_edge_shapes(d, B):
    """d not a multiple of 8, batch 1, odd sizes - through the real kernels (and, in the sanitizer runs,
    under ASan / TSan)."""
    Z = staged(R.KIND_BINARY, d, B, seed=d * 100 + B)
    w0 = 0.1 * torch.randn(d + 1)

    def run():
        outs = []
        w = w0.clone()
        st = torch.zeros(4, dtype=torch.float64)
        sc = F.linear_minibatch_scratch(d, DEV)
        F.linear_minibatch(Z, d, w, st, sc, F.KIND_PA, 1, 0.5, 0.05, 0.01)
        outs += [w, st]
        w2 = w0.clone()
        st2 = torch.zeros(4, dtype=torch.float64)
        F.linear_exact(Z, d, w2, st2, F.KIND_PA, 1, 0.5, 0.05, 0.01)
        outs += [w2, st2, F.linear_predict(Z, d, w0)]
        ring = torch.zeros(1, Z.shape[1])
        Zh = Z.clone()
        F.holdout_swap(Zh, ring, d, 7)                      # capacity 1, count starting mid-decade
        outs += [Zh, ring]
        ss = torch.zeros(1 + 2 * d, dtype=torch.float64)
        sx = torch.zeros(3 + 4 * d, dtype=torch.float64)
        Zs = Z.clone()
        F.standard_scaler(Zs, d, ss, sx, True)
        outs += [Zs, ss]
        stq = torch.zeros(4, dtype=torch.float64)
        F.sq_residual(Z, d, torch.cat([w0, -torch.ones(1)]), stq)
        outs += [stq, F.row_sqnorm(Z, d + 1)]
        return outs
    ref, nat = both(run)
    for i, (r, n) in enumerate(zip(ref, nat)):
        assert torch.allclose(r.double(), n.double(), rtol=2e-3, atol=2e-4, equal_nan=True), (i, d, B)


@pytest.mark.parametrize("classes", [0, 4])
def test_nn_glue_kernels(classes):
    """csrc/nn.cu: activation, output loss / gradient (soft-max cross entropy and squared error), relu mask, SGD step."""
    B, d, od = 37, 9, (classes if classes > 1 else 1)
    Z = staged(R.KIND_MULTICLASS if classes > 1 else R.KIND_REGRESSION, d, B, seed=3, n_aux=max(classes, 1))
    Z[5, d] = 0.0                                            # an invalid row contributes nothing
    H0 = torch.randn(B, F.pad32(od + 6))
    out0 = torch.randn(B, F.pad32(od + 1))
    W0, gW0 = torch.randn(6, 32), torch.randn(6, 32)

    def run():
        H = H0.clone()
        F.nn_hidden_act(H, od + 3)
        G = torch.full((B, F.pad32(od)), 7.0)
        st = torch.zeros(4, dtype=torch.float64)
        nv = torch.zeros(1)
        F.nn_out_grad(out0, Z, d, od, classes, G, st, nv)
        g = torch.randn(B, 32, generator=torch.Generator().manual_seed(1))
        F.nn_relu_mask(g, H0, 20)
        W = W0.clone()
        F.nn_sgd_step(W, gW0, 0.1, nv)
        return H, G, st, nv, g, W
    ref, nat = both(run)
    for i, (r, n) in enumerate(zip(ref, nat)):
        assert torch.allclose(r.double(), n.double(), rtol=1e-4, atol=1e-5), i
    assert float(nat[3]) == B - 1 and float(nat[1][5].abs().sum()) == 0.0


def test_engine_nn_on_emulated_kernels():
    d, B = 10, 64
    batches = [R.generate_batch(R.KIND_REGRESSION, d, B, seed=13, batch_index=i) for i in range(5)]
    out = []
    for impl in ("torch", "emu"):
        with F.use_impl(impl):
            e = _mk_engine({"name": "NN", "hyperParameters": {"hiddenLayers": [8, 5], "learningRate": 0.05}}, d=d)
            for b in batches:
                e.train_batch(b.clone(), d)
            l = e.pipelines[1].learner
            out.append((l.buf.clone(), l.counters(), l.predict(batches[0].clone()).clone()))
    (bt, ct, pt), (be, ce, pe) = out
    assert ct["fitted"] == ce["fitted"] > 0 and abs(ct["loss_sum"] - ce["loss_sum"]) <= 1e-3 * abs(ct["loss_sum"])
    assert torch.allclose(bt, be, rtol=1e-3, atol=1e-5) and torch.allclose(pt, pe, rtol=1e-3, atol=1e-4)
