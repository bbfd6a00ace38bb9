"""GPU learners pinned to the fp64 oracle, end to end through the engine: ORR on fp16 rows, K-means and MultiClassPA
(mini-batch rules of oracle/learners.py), the way tests/test_engine_gpu.py pins PA and ORR on fp32 rows."""
import json

import numpy as np
import pytest
import torch

from omldm_b200.engine import Engine
from omldm_b200.ops import functional as F
from omldm_b200.oracle.learners import make_oracle_learner
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters

pytestmark = pytest.mark.gpu


def _engine(name, hyper, d, tc=None):
    eng = Engine(JobParameters(parallelism=1, testSetSize=0), device="cuda")
    eng.handle_request({"id": 1, "request": "Create", "learner": {"name": name, "hyperParameters": hyper},
                        "trainingConfiguration": dict({"protocol": "Synchronous", "nFeatures": d}, **(tc or {}))})
    return eng


def test_orr_on_fp16_rows_matches_the_oracle_on_the_same_rounded_records():
    """fp16 staged rows through the CTA-pair Gram kernel, the fp16 loss pass and the CG / Cholesky solvers: the oracle
    is fed the records as fp16 rounds them, so what remains is accumulation and solver error."""
    d, B, steps = 200, 1024, 8
    eng = _engine("ORR", {"lambda": 1.0, "solveEvery": 2}, d)
    o = make_oracle_learner("ORR", {"lambda": 1.0})
    for i in range(steps):
        Zh = R.generate_batch(R.KIND_REGRESSION, d, B, seed=21, batch_index=i, wire="f16s")
        Z = F.expand_rows_f16s(Zh, d)
        o.fit_batch(Z[:, :d].double().numpy(), Z[:, d + 1].double().numpy())
        eng.train_batch(Zh.cuda(), d)
    out = [json.loads(x) for x in eng.handle_request({"id": 1, "request": "Query", "requestId": 1})]
    from omldm_b200.api.bucketing import reassemble
    from omldm_b200.api.pojos import QueryResponse
    full = reassemble([QueryResponse.from_dict(r) for r in out])
    w = np.array(full.learner.parameters["weights"])
    assert np.allclose(w, o.w, rtol=1e-3, atol=1e-4), float(np.abs(w - o.w).max())
    A = np.array(full.learner.parameters["a"]).reshape(d + 1, d + 1)
    Aref = o.G[:d + 1, :d + 1] + np.eye(d + 1)
    assert np.allclose(A, Aref, rtol=0, atol=2e-5 * np.abs(Aref).max())
    assert full.dataFitted == B * steps
    eng.close()


def test_kmeans_centroids_match_the_oracle():
    d, K, B, steps = 32, 16, 2048, 8
    eng = _engine("K-means", {"k": K}, d)
    o = make_oracle_learner("K-means", {"k": K})
    nb = max(2 * K, 64)      # the engine's bootstrap: the first nb records in exact (MacQueen) order, then mini-batches
    for i in range(steps):
        Z = R.generate_batch(R.KIND_CLUSTERS, d, B, seed=11, batch_index=i, n_aux=K)
        X = Z[:, :d].double().numpy()
        if i == 0:
            for x in X[:nb]:
                o.fit(x, 0.0)
            o.fit_batch(X[nb:])
        else:
            o.fit_batch(X)
        eng.train_batch(Z.cuda(), d)
    r = json.loads(eng.handle_request({"id": 1, "request": "Query", "requestId": 1})[0])
    p = r["learner"]["parameters"]
    C = np.array(p["centroids"]).reshape(K, d)
    n = np.array(p["counts"])
    assert n.sum() == B * steps
    assert np.abs(n - o.S[:, -1]).sum() <= 0.005 * B * steps        # tf32 distances may flip a near tie
    # a record within tf32 rounding of two centroids may go either way and moves a small cluster's centroid visibly:
    # tight on the member-weighted average, loose on the single worst coordinate
    err = np.abs(C - o.centroids()).max(1)
    assert float((err * o.S[:, -1]).sum() / o.S[:, -1].sum()) < 3e-2, err
    assert float(err.max()) < 0.5, err
    assert abs(r["cumulativeLoss"] - o.cum_loss) <= 2e-2 * o.cum_loss
    eng.close()


@pytest.mark.parametrize("sparse", [False, True])
def test_multiclass_pa_weights_match_the_oracle(sparse):
    """Dense path (two tcgen05 GEMMs + coefficient kernel) and the sparse-direct path (gathers / scatters from the
    non-zeros) against the oracle's mini-batch rule on the same records."""
    C = 6
    hyper = {"nClasses": C, "C": 1.0}
    if not sparse:
        d, B, steps = 64, 512, 20
        eng = _engine("MultiClassPA", hyper, d)
        o = make_oracle_learner("MultiClassPA", hyper)
        for i in range(steps):
            Z = R.generate_batch(R.KIND_MULTICLASS, d, B, seed=3, batch_index=i, n_aux=C)
            o.fit_batch(Z[:, :d].double().numpy(), Z[:, d + 1].double().numpy())
            eng.train_batch(Z.cuda(), d)
    else:
        d, B, nnz, steps = 2000, 256, 32, 12
        jp = JobParameters(parallelism=1, testSetSize=0)
        jp.extra["sparseDirect"] = "true"
        eng = Engine(jp, device="cuda")
        eng.handle_request({"id": 1, "request": "Create", "learner": {"name": "MultiClassPA", "hyperParameters": hyper},
                            "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
        o = make_oracle_learner("MultiClassPA", hyper)
        g = torch.Generator().manual_seed(5)
        planes = torch.randn(C, d, generator=g)
        for i in range(steps):
            # unique column indices per record: with duplicates the dense rule's step size sees (sum v)^2 and the
            # sparse kernel sum v^2 (documented in include/omldm_kernels.h)
            idx = torch.stack([torch.randperm(d, generator=g)[:nnz] for _ in range(B)]).int()
            val = torch.randn(B, nnz, generator=g)
            X = torch.zeros(B, d, dtype=torch.float64)
            X.scatter_(1, idx.long(), val.double())
            tgt = (X.float() @ planes.t()).argmax(1).float()
            slot = torch.cat([idx.reshape(-1).view(torch.uint8), val.reshape(-1).view(torch.uint8), tgt.view(torch.uint8)])
            o.fit_batch(X.numpy(), tgt.double().numpy())
            eng.train_sparse(slot.cuda(), B, nnz, d)
    r = [json.loads(x) for x in eng.handle_request({"id": 1, "request": "Query", "requestId": 1})]
    from omldm_b200.api.bucketing import reassemble
    from omldm_b200.api.pojos import QueryResponse
    full = reassemble([QueryResponse.from_dict(x) for x in r])
    W = np.array(full.learner.parameters["a"]).reshape(C, -1)
    Wo = np.asarray(o.W)[:, : W.shape[1]]
    rel = np.linalg.norm(W - Wo) / np.linalg.norm(Wo)
    assert rel < 5e-2, rel
    assert full.dataFitted == B * steps
    assert abs(full.cumulativeLoss - o.cum_loss) <= 3e-2 * max(1.0, o.cum_loss)
    eng.close()
