"""Conjugate-gradient weight refresh of online ridge regression (csrc/ridge_cg.cu and its torch restatement):
agreement with a direct fp64 solve, warm-start behaviour on a stream, and the fall-back to Cholesky."""
import pytest
import torch

from omldm_b200.engine import Engine
from omldm_b200.ops import functional as F
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters



def _system(n, m, seed, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    X = torch.randn(m, n, generator=g) * (1.0 + 3.0 * torch.rand(n, generator=g))     # uneven feature scales
    y = X @ torch.randn(n, generator=g) + 0.1 * torch.randn(m, generator=g)
    ld = F.pad32(n + 1)
    W = torch.zeros(ld, ld)
    W[:n, :n] = X.t() @ X
    W[:n, n] = X.t() @ y
    return W.to(device)


@pytest.mark.parametrize("n,m", [(5, 50), (64, 2000), (257, 5000)])
def test_cg_matches_direct_solve(n, m):
    W = _system(n, m, n)
    x = torch.zeros(n)
    work, scal, info = F.ridge_cg_workspace(n, "cpu", 64)
    F.ridge_cg(W, n, 1.0, x, work, scal, info, max_iter=64, tol=1e-6)
    ref = torch.linalg.solve(W[:n, :n].double() + torch.eye(n, dtype=torch.float64), W[:n, n].double()).float()
    assert float(info[1]) <= 1e-6 and 0 < int(info[0]) <= 64
    assert torch.allclose(x, ref, rtol=2e-3, atol=2e-4)
    # a converged start needs no iteration at all
    F.ridge_cg(W, n, 1.0, x, work, scal, info, max_iter=64, tol=1e-4)
    assert int(info[0]) == 0


def test_cg_reports_non_convergence():
    n = 40
    W = _system(n, 30, 1)                       # rank deficient, tiny lambda: cannot converge in 3 iterations
    x = torch.zeros(n)
    work, scal, info = F.ridge_cg_workspace(n, "cpu", 3)
    F.ridge_cg(W, n, 1e-6, x, work, scal, info, max_iter=3, tol=1e-7)
    assert int(info[0]) == 3 and float(info[1]) > 1e-7


def _orr(solver, d, every=2, extra=None):
    e = Engine(JobParameters(parallelism=1, testSetSize=32), device="cpu")
    hp = {"lambda": 1.0, "solver": solver, "solveEvery": every}
    hp.update(extra or {})
    e.handle_request({"id": 1, "request": "Create", "learner": {"name": "ORR", "hyperParameters": hp},
                      "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
    return e


def test_orr_with_cg_refresh_tracks_cholesky_and_warm_start_pays():
    d = 48
    a, b = _orr("cholesky", d), _orr("cg", d)
    iters = []
    for i in range(24):
        Z = R.generate_batch(R.KIND_REGRESSION, d, 256, seed=3, batch_index=i)
        a.train_batch(Z.clone(), d)
        b.train_batch(Z.clone(), d)
        iters.append(b.pipelines[1].learner.cg_last[0])
    la, lb = a.pipelines[1].learner, b.pipelines[1].learner
    la.solve()
    lb.solve()
    assert lb.cg_failures == 0 and lb.solver == "cg"
    assert torch.allclose(la.wt, lb.wt, rtol=1e-3, atol=1e-4)
    ca, cb = la.counters(), lb.counters()
    assert ca["fitted"] == cb["fitted"] and abs(ca["loss_sum"] - cb["loss_sum"]) <= 1e-3 * ca["loss_sum"]
    late = [k for k in iters[-8:] if k > 0]
    assert late and max(late) <= 4, f"warm-started refreshes should need a handful of iterations, got {iters}"


def test_orr_cg_falls_back_to_direct_solve():
    d = 30
    e = _orr("cg", d, every=1, extra={"cgMaxIter": 1, "cgTol": 1e-9, "lambda": 1e-4})
    for i in range(4):
        e.train_batch(R.generate_batch(R.KIND_REGRESSION, d, 64, seed=5, batch_index=i), d)
    l = e.pipelines[1].learner
    assert l.cg_failures >= 1                   # one iteration cannot reach 1e-9: every refresh was redone directly
    ref = _orr("cholesky", d, every=1, extra={"lambda": 1e-4})
    for i in range(4):
        ref.train_batch(R.generate_batch(R.KIND_REGRESSION, d, 64, seed=5, batch_index=i), d)
    assert torch.allclose(l.wt, ref.pipelines[1].learner.wt, rtol=1e-4, atol=1e-5)


@pytest.mark.gpu
def test_ridge_cg_kernel_matches_fp64_solve():
    for n, m in ((7, 100), (129, 3000), (1001, 20000)):
        W = _system(n, m, n + 1, "cuda")
        x = torch.zeros(n, device="cuda")
        work, scal, info = F.ridge_cg_workspace(n, "cuda", 64)
        F.ridge_cg(W, n, 1.0, x, work, scal, info, max_iter=64, tol=1e-6)
        torch.cuda.synchronize()
        ref = torch.linalg.solve(W[:n, :n].double().cpu() + torch.eye(n, dtype=torch.float64),
                                 W[:n, n].double().cpu()).float()
        assert float(info[1]) <= 1e-5, (n, info.tolist())
        assert torch.allclose(x.cpu(), ref, rtol=5e-3, atol=5e-4), n
        F.ridge_cg(W, n, 1.0, x, work, scal, info, max_iter=64, tol=1e-4)      # warm start: already converged
        torch.cuda.synchronize()
        assert int(info[0]) == 0


@pytest.mark.gpu
def test_orr_gpu_cg_refresh_matches_cholesky():
    d = 200
    outs = []
    for solver in ("cholesky", "cg"):
        e = Engine(JobParameters(parallelism=1, testSetSize=64))
        e.handle_request({"id": 1, "request": "Create",
                          "learner": {"name": "ORR", "hyperParameters": {"lambda": 1.0, "solver": solver, "solveEvery": 4}},
                          "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
        for i in range(20):
            e.train_batch(R.generate_batch(R.KIND_REGRESSION, d, 1024, seed=7, batch_index=i).cuda(), d)
        l = e.pipelines[1].learner
        l.solve()
        torch.cuda.synchronize()
        outs.append((l.wt.cpu().clone(), l.counters(), l.cg_failures))
        e.close()
    assert outs[1][2] == 0
    assert torch.allclose(outs[0][0], outs[1][0], rtol=2e-3, atol=2e-4)
    assert outs[0][1]["fitted"] == outs[1][1]["fitted"]
