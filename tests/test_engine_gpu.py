"""GPU engine (native kernels) against the fp64 oracle learners, end to end through JSON requests."""
import json

import numpy as np
import pytest
import torch

from omldm_b200.engine import Engine
from omldm_b200.ops import _lib
from omldm_b200.oracle.learners import make_oracle_learner
from omldm_b200.oracle.preprocessors import PolynomialFeatures, StandardScaler
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters

pytestmark = pytest.mark.gpu


def run_engine(name, hyper, kind, d, B, steps, pre=(), n_aux=0, proto="Synchronous", device="cuda"):
    eng = Engine(JobParameters(parallelism=1, testSetSize=0), device=device)
    eng.handle_request({"id": 1, "request": "Create", "learner": {"name": name, "hyperParameters": hyper},
                        "preProcessors": list(pre), "trainingConfiguration": {"protocol": proto, "nFeatures": d}})
    batches = [R.generate_batch(kind, d, B, seed=11, batch_index=i, n_aux=n_aux) for i in range(steps)]
    for b in batches:
        eng.train_batch(b.to(device), d)
    out = eng.handle_request({"id": 1, "request": "Query", "requestId": 5})
    return eng, batches, [json.loads(o) for o in out]


def oracle_fit(name, hyper, batches, d, pre=()):
    o = make_oracle_learner(name, hyper)
    for b in batches:
        X, y = b[:, :d].double().numpy(), b[:, d + 1].double().numpy()
        for p in pre:
            X = p.fit_transform_batch(X)
        o.fit_batch(X, y)
    return o


@pytest.mark.parametrize("name,kind,hyper", [
    ("PA", R.KIND_BINARY, {"C": 0.1}), ("RegressorPA", R.KIND_REGRESSION, {"C": 0.5, "epsilon": 0.01}),
    ("SVM", R.KIND_BINARY, {"lambda": 0.01})])
def test_linear_minibatch_matches_oracle(name, kind, hyper):
    d = 28
    eng, batches, resp = run_engine(name, dict(hyper, trainingMode="minibatch"), kind, d, 256, 20)
    o = oracle_fit(name, hyper, batches, d)
    a = np.array(resp[0]["learner"]["parameters"]["a"])
    assert np.allclose(a, o.w[:-1], rtol=5e-3, atol=5e-4)
    assert resp[0]["dataFitted"] == 256 * 20
    assert abs(resp[0]["cumulativeLoss"] - o.cum_loss) <= 1e-3 * max(1.0, o.cum_loss)
    assert _lib.launch_count() > 0
    eng.close()


def test_orr_matches_oracle_and_buckets():
    d = 200
    eng, batches, resp = run_engine("ORR", {"lambda": 1.0, "solveEvery": 1}, R.KIND_REGRESSION, d, 512, 6)
    o = oracle_fit("ORR", {"lambda": 1.0}, batches, d)
    assert len(resp) > 2   # (d+1)^2 parameters need several 10 000-element buckets
    from omldm_b200.api.bucketing import reassemble
    from omldm_b200.api.pojos import QueryResponse
    full = reassemble([QueryResponse.from_dict(r) for r in resp])
    w = np.array(full.learner.parameters["weights"])
    assert np.allclose(w, o.w, rtol=2e-2, atol=2e-3)      # tf32 Gram
    A = np.array(full.learner.parameters["a"]).reshape(d + 1, d + 1)
    Aref = o.G[:d + 1, :d + 1] + np.eye(d + 1)
    assert np.allclose(A, Aref, rtol=0, atol=2e-3 * np.abs(Aref).max())
    assert full.dataFitted == 512 * 6
    eng.close()


def test_fgm_config_pipeline_runs_and_learns():
    """cfg 3: PolynomialFeatures + StandardScaler + RegressorPA."""
    d = 8
    pre = [{"name": "PolynomialFeatures", "hyperParameters": {"degree": 2}}, {"name": "StandardScaler"}]
    hyper = {"C": 0.5, "epsilon": 0.01, "trainingMode": "minibatch"}
    eng, batches, resp = run_engine("RegressorPA", hyper, R.KIND_REGRESSION, d, 512, 30, pre=pre, proto="FGM")
    o = oracle_fit("RegressorPA", hyper, batches, d, pre=[PolynomialFeatures({"degree": 2}), StandardScaler()])
    a = np.array(resp[0]["learner"]["parameters"]["a"])
    assert a.shape[0] == 44
    assert np.allclose(a, o.w[:-1], rtol=2e-2, atol=2e-3)
    eng.close()


def test_kmeans_and_mcpa_learn():
    d, K = 32, 16
    eng, batches, resp = run_engine("K-means", {"k": K}, R.KIND_CLUSTERS, d, 2048, 8, n_aux=K)
    p = resp[0]["learner"]["parameters"]
    assert sum(p["counts"]) == pytest.approx(2048 * 8, rel=1e-6)
    eng.close()
    d, C = 64, 6
    eng = Engine(JobParameters(parallelism=1), device="cuda")
    eng.handle_request({"id": 1, "request": "Create",
                        "learner": {"name": "MultiClassPA", "hyperParameters": {"nClasses": C, "C": 1.0}},
                        "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
    for i in range(40):
        eng.train_batch(R.generate_batch(R.KIND_MULTICLASS, d, 512, seed=3, batch_index=i, n_aux=C).cuda(), d)
    r = json.loads(eng.handle_request({"id": 1, "request": "Query", "requestId": 1})[0])
    assert r["score"] > 0.7
    eng.close()


def test_e2e_host_path_and_forecast():
    d = 28
    eng = Engine(JobParameters(parallelism=1), device="cuda")
    eng.handle_request({"id": 3, "request": "Create", "learner": {"name": "PA", "hyperParameters": {"C": 0.1}},
                        "trainingConfiguration": {"nFeatures": d}})
    topic = R.Topic(d, 1024, n_slots=4)
    topic.start_synthetic(R.KIND_BINARY, seed=5, max_batches=12)
    n = 0
    while True:
        got = topic.get(5000)
        if got is None:
            break
        slot, Zh, nr = got
        loss = eng.train_batch_host(Zh, d)
        torch.cuda.synchronize()
        topic.release(slot)
        n += nr
    assert n == 12 * 1024 and float(loss[2]) > 0
    preds = eng.forecast_batch(R.generate_batch(R.KIND_BINARY, d, 8, seed=5, batch_index=99).cuda())
    assert len(preds) == 8 and "prediction" in preds[0]
    js = eng.terminate()
    assert js.statistics[0].fitted > 0 and js.statistics[0].score > 0.8
    topic.stop()
    eng.close()


def test_multi_pipeline_exact_batched_launch_matches_separate_pipelines():
    """Many tiny pipelines share one kernel launch (one warp each) and still equal the oracle."""
    d, B = 28, 256
    specs = [("PA", {"C": 0.1}), ("PA", {"C": 1.0, "updateType": "PA-II"}), ("SVM", {"lambda": 0.01}),
             ("RegressorPA", {"C": 0.5, "epsilon": 0.05})]
    eng = Engine(JobParameters(parallelism=1, testSetSize=0, test=False), device="cuda")   # no curve samples: count launches
    for i, (name, hyper) in enumerate(specs):
        eng.handle_request({"id": i + 1, "request": "Create",
                            "learner": {"name": name, "hyperParameters": dict(hyper, trainingMode="exact")},
                            "trainingConfiguration": {"protocol": "Asynchronous", "nFeatures": d}})
    batches = [R.generate_batch(R.KIND_BINARY, d, B, seed=21, batch_index=i) for i in range(6)]
    before = _lib.launch_count()
    for b in batches:
        eng.train_batch(b.cuda(), d)
    torch.cuda.synchronize()
    assert _lib.launch_count() - before <= len(batches) * 2      # one shared launch per batch (+ nothing per pipeline)
    for i, (name, hyper) in enumerate(specs):
        o = make_oracle_learner(name, hyper)
        for b in batches:
            for r in b:
                o.fit(r[:d].double().numpy(), float(r[d + 1]))
        resp = json.loads(eng.handle_request({"id": i + 1, "request": "Query", "requestId": 1})[0])
        assert np.allclose(resp["learner"]["parameters"]["a"], o.w[:-1], rtol=2e-3, atol=2e-4), name
        assert resp["dataFitted"] == B * len(batches)
    eng.close()


@pytest.mark.parametrize("name,hyper,kind,wire,n_bufs", [
    ("ORR", {"lambda": 1.0, "solveEvery": 4}, R.KIND_REGRESSION, "f16s", 3),
    ("ORR", {"lambda": 1.0, "solveEvery": 4}, R.KIND_REGRESSION, "f16s", 1),
    ("PA", {"C": 0.1, "trainingMode": "minibatch"}, R.KIND_BINARY, "f32", 2),
])
def test_pipelined_holdout_gives_the_same_model_and_ring(name, hyper, kind, wire, n_bufs):
    """`pipelinedHoldout`: the holdout exchange of batch i + 1 runs on its own stream behind the training kernels of
    batch i.  Same weights, statistics, holdout ring and record count as the default in-graph order - with distinct
    buffers per step and with ONE buffer refilled for every step (the exchange must then wait for the previous step)."""
    d, B, steps = 60, 4096, 14

    def run(pipelined):
        jp = JobParameters(parallelism=1, testSetSize=64)
        jp.extra["pipelinedHoldout"] = "true" if pipelined else "false"
        eng = Engine(jp, device="cuda")
        eng.handle_request({"id": 1, "request": "Create", "learner": {"name": name, "hyperParameters": hyper},
                            "preProcessors": [], "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
        bufs = [None] * n_bufs
        for i in range(steps):
            src = R.generate_batch(kind, d, B, seed=3, batch_index=i, wire=wire).cuda()
            k = i % n_bufs
            if bufs[k] is None:
                bufs[k] = torch.empty_like(src)
            eng._buffer_free(bufs[k])        # what an owner of a reused buffer does in pipelined mode
            bufs[k].copy_(src)
            eng.train_batch(bufs[k], d)
        out = [json.loads(o) for o in eng.handle_request({"id": 1, "request": "Query", "requestId": 5})]
        lrn = eng.pipelines[1].learner
        torch.cuda.synchronize()
        res = (lrn.stats.clone(), eng.holdout.clone(), eng.count, out, any(isinstance(g, tuple) for g in eng._graphs.values()))
        eng.close()
        return res

    s0, h0, c0, o0, piped0 = run(False)
    s1, h1, c1, o1, piped1 = run(True)
    assert piped1 and not piped0                      # the pipelined run really replayed split graphs
    assert c0 == c1 and torch.equal(h0, h1)
    assert torch.allclose(s0, s1, rtol=1e-6)
    p0, p1 = o0[0]["learner"]["parameters"], o1[0]["learner"]["parameters"]
    assert np.allclose(np.asarray(p0["a"]), np.asarray(p1["a"]), rtol=1e-4, atol=1e-6)
