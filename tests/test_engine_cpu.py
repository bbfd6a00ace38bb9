"""GPU-engine code path on CPU tensors (torch implementation of every op) against the fp64 oracle,
plus the native host library (rings, producers, JSON scanner)."""
import json

import numpy as np
import pytest
import torch

from omldm_b200.engine import Engine
from omldm_b200.ops import functional as F
from omldm_b200.oracle.learners import make_oracle_learner
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters


def run(name, hyper, kind, d, B, steps, n_aux=0, pre=(), holdout=0):
    eng = Engine(JobParameters(parallelism=1, testSetSize=holdout), device="cpu")
    eng.handle_request({"id": 1, "request": "Create", "learner": {"name": name, "hyperParameters": hyper},
                        "preProcessors": list(pre), "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
    batches = [R.generate_batch(kind, d, B, seed=4, batch_index=i, n_aux=n_aux) for i in range(steps)]
    for b in batches:
        eng.train_batch(b.clone(), d)
    resp = [json.loads(o) for o in eng.handle_request({"id": 1, "request": "Query", "requestId": 2})]
    return eng, batches, resp


def oracle(name, hyper, batches, d):
    o = make_oracle_learner(name, hyper)
    for b in batches:
        o.fit_batch(b[:, :d].double().numpy(), b[:, d + 1].double().numpy())
    return o


@pytest.mark.parametrize("name,kind,hyper", [("PA", R.KIND_BINARY, {"C": 0.1}),
                                             ("RegressorPA", R.KIND_REGRESSION, {"C": 0.5, "epsilon": 0.01}),
                                             ("SVM", R.KIND_BINARY, {"lambda": 0.01})])
def test_linear_minibatch_equals_oracle(name, kind, hyper):
    d = 20
    eng, batches, resp = run(name, dict(hyper, trainingMode="minibatch"), kind, d, 128, 10)
    o = oracle(name, hyper, batches, d)
    assert np.allclose(resp[0]["learner"]["parameters"]["a"], o.w[:-1], rtol=1e-3, atol=1e-4)
    assert resp[0]["protocol"] == "CentralizedTraining" and resp[0]["dataFitted"] == 1280


def test_linear_exact_equals_oracle_per_record():
    d = 10
    eng, batches, resp = run("PA", {"C": 0.1, "trainingMode": "exact"}, R.KIND_BINARY, d, 64, 3)
    o = make_oracle_learner("PA", {"C": 0.1})
    for b in batches:
        for r in b:
            o.fit(r[:d].double().numpy(), float(r[d + 1]))
    assert np.allclose(resp[0]["learner"]["parameters"]["a"], o.w[:-1], rtol=1e-4, atol=1e-5)


def test_orr_kmeans_mcpa_nn_against_oracle():
    d = 16
    eng, batches, resp = run("ORR", {"lambda": 1.0, "solveEvery": 1}, R.KIND_REGRESSION, d, 256, 4)
    o = oracle("ORR", {"lambda": 1.0}, batches, d)
    assert np.allclose(resp[0]["learner"]["parameters"]["weights"], o.w, rtol=1e-3, atol=1e-4)
    eng, batches, resp = run("K-means", {"k": 6}, R.KIND_CLUSTERS, d, 512, 4, n_aux=6)
    assert abs(sum(resp[0]["learner"]["parameters"]["counts"]) - 2048) < 1e-3
    eng, batches, resp = run("MultiClassPA", {"nClasses": 4, "C": 1.0}, R.KIND_MULTICLASS, d, 256, 20, n_aux=4)
    o = oracle("MultiClassPA", {"nClasses": 4, "C": 1.0}, batches, d)
    W = np.array(resp[0]["learner"]["parameters"]["a"]).reshape(4, d)
    assert np.allclose(W, o.W[:, :-1], rtol=5e-3, atol=5e-4)
    eng, batches, resp = run("NN", {"hiddenLayers": [8], "learningRate": 0.05}, R.KIND_REGRESSION, d, 64, 10)
    o = oracle("NN", {"hiddenLayers": [8], "learningRate": 0.05}, batches, d)
    assert np.allclose(np.array(resp[0]["learner"]["parameters"]["layer0"]), o.layers[0].reshape(-1), rtol=1e-3, atol=1e-4)


def test_holdout_rule_matches_reference_semantics_and_scores():
    d = 8
    eng, batches, resp = run("PA", {"C": 0.1}, R.KIND_BINARY, d, 500, 6, holdout=64)
    # 3000 records: 600 holdouts, 64 kept, the rest (evicted) trained
    assert resp[0]["dataFitted"] == 3000 - 64 and resp[0]["score"] > 0.8
    assert eng._holdout_rows().shape[0] == 64


def test_preprocessor_pipeline_and_warm_start():
    d = 6
    pre = [{"name": "PolynomialFeatures", "hyperParameters": {"degree": 2}}, {"name": "MinMaxScaler"}]
    eng, batches, resp = run("RegressorPA", {"C": 0.5}, R.KIND_REGRESSION, d, 128, 5, pre=pre)
    assert len(resp[0]["learner"]["parameters"]["a"]) == 27
    assert [p["name"] for p in resp[0]["preprocessors"]] == ["PolynomialFeatures", "MinMaxScaler"]
    params = resp[0]["learner"]["parameters"]
    eng2 = Engine(JobParameters(parallelism=1), device="cpu")
    eng2.handle_request({"id": 9, "request": "Create", "learner": {"name": "RegressorPA", "parameters": params},
                         "preProcessors": pre, "trainingConfiguration": {"nFeatures": d}})
    r2 = json.loads(eng2.handle_request({"id": 9, "request": "Query", "requestId": 1})[0])
    assert np.allclose(r2["learner"]["parameters"]["a"], params["a"], atol=1e-6)


def test_checkpoint_roundtrip_and_delete():
    d = 12
    eng, batches, resp = run("PA", {"C": 0.1}, R.KIND_BINARY, d, 128, 4, holdout=32)
    sd = eng.state_dict()
    eng2 = Engine(JobParameters(parallelism=1, testSetSize=32), device="cpu")
    eng2.load_state_dict(sd)
    r2 = json.loads(eng2.handle_request({"id": 1, "request": "Query", "requestId": 5})[0])
    assert np.allclose(r2["learner"]["parameters"]["a"], resp[0]["learner"]["parameters"]["a"])
    assert r2["dataFitted"] == resp[0]["dataFitted"]
    eng2.handle_request({"id": 1, "request": "Delete"})
    assert eng2.handle_request({"id": 1, "request": "Query", "requestId": 6}) == []
    js = eng.terminate()
    assert js.statistics[0].fitted == resp[0]["dataFitted"]


def test_topic_ring_producer_and_json_scanner():
    t = R.Topic(5, 64, n_slots=3, pin=False)
    t.start_synthetic(R.KIND_BINARY, seed=2, max_batches=7, n_threads=2)
    n, seen = 0, []
    while True:
        g = t.get(5000)
        if g is None:
            break
        slot, Z, nr = g
        assert bool((Z[:, 5] == 1).all()) and set(Z[:, 6].unique().tolist()) <= {-1.0, 1.0}
        n += nr
        t.release(slot)
    assert n == 7 * 64 and t.consumed == n
    t.stop()
    t2 = R.Topic(3, 4, n_slots=2, pin=False)
    rows, nf, used = R.parse_json_records(
        b'{"numericalFeatures":[1.5,2,3],"target":1.0,"operation":"training"}\nEOS\n{"numericalFeatures":[1,2]}\n'
        b'{"numericalFeatures":[4,5],"discreteFeatures":[6],"operation":"forecasting"}\n', 3, 10)
    assert rows.shape[0] == 2 and nf == 1 and rows[0, :5].tolist() == [1.5, 2.0, 3.0, 1.0, 1.0]
    assert rows[1, :4].tolist() == [4.0, 5.0, 6.0, 2.0]
    assert t2.put(rows) and t2.get(100)[2] == 2
    same = R.generate_batch(R.KIND_REGRESSION, 7, 16, seed=5, batch_index=3)
    assert torch.equal(same, R.generate_batch(R.KIND_REGRESSION, 7, 16, seed=5, batch_index=3))


def test_hoeffding_tree_device_statistics_path_learns():
    """HT with leaf statistics accumulated by the (torch-implemented on CPU) tree kernels."""
    d = 10
    eng = Engine(JobParameters(parallelism=1, testSetSize=64), device="cpu")
    eng.handle_request({"id": 1, "request": "Create",
                        "learner": {"name": "HT", "hyperParameters": {"nClasses": 2, "gracePeriod": 200}},
                        "trainingConfiguration": {"nFeatures": d}})
    rng = np.random.default_rng(0)
    for i in range(30):
        X = rng.standard_normal((256, d)).astype(np.float32)
        y = np.where((X[:, 2] > 0.3) ^ (X[:, 5] < -0.5), 1.0, -1.0)
        eng.train_batch(F.stage_rows(torch.from_numpy(np.hstack([X, y[:, None]]).astype(np.float32)), d), d)
    r = json.loads(eng.handle_request({"id": 1, "request": "Query", "requestId": 1})[0])
    assert r["protocol"] == "SingleLearner" and r["score"] > 0.9
    ds = r["learner"]["dataStructure"]
    assert len(ds["feature"]) >= 5 and set(f for f in ds["feature"] if f >= 0) <= {2, 5}
    sd = eng.state_dict()
    eng2 = Engine(JobParameters(parallelism=1, testSetSize=64), device="cpu")
    eng2.load_state_dict(sd)
    r2 = json.loads(eng2.handle_request({"id": 1, "request": "Query", "requestId": 2})[0])
    assert r2["learner"]["dataStructure"]["feature"] == ds["feature"]


def test_records_before_the_first_create_are_buffered_and_trained_on():
    """FlinkSpoke.scala:69-80 / SpokeLogic.scala:32: without a pipeline, training records wait in a bounded FIFO
    (100 000 records) and are trained on, oldest first, as soon as a pipeline exists."""
    import torch
    from omldm_b200.engine import Engine
    from omldm_b200.stream import ring as R
    from omldm_b200.utils.config import RECORD_BUFFER_CAP, JobParameters
    d = 6
    a = Engine(JobParameters(parallelism=1, testSetSize=16), device="cpu")
    b = Engine(JobParameters(parallelism=1, testSetSize=16), device="cpu")
    create = {"id": 1, "request": "Create", "learner": {"name": "PA", "hyperParameters": {"C": 0.1}},
              "trainingConfiguration": {"protocol": "Synchronous"}}
    batches = [R.generate_batch(R.KIND_BINARY, d, 100, seed=2, batch_index=i) for i in range(6)]
    b.handle_request(create)                            # b: pipeline first, then the whole stream
    for i, Z in enumerate(batches):
        if i == 3:
            assert a._buffered_rows == 300 and a.count == 0 and not a.pipelines
            a.handle_request(create)                    # a: three batches arrive before the Create
        a.train_batch(Z.clone(), d)
        b.train_batch(Z.clone(), d)
    assert a._buffered_rows == 0 and a.count == b.count == 600
    assert torch.equal(a.pipelines[1].learner.w, b.pipelines[1].learner.w)          # same records, same order
    # the buffer is bounded: the oldest records are dropped
    c = Engine(JobParameters(parallelism=1), device="cpu")
    big = torch.zeros(60000, 32)
    big[:, d] = 1.0
    c.train_batch(big.clone(), d)
    c.train_batch(big.clone(), d)
    assert c._buffered_rows == RECORD_BUFFER_CAP and c.dropped_records == 20000


def test_create_is_parked_until_every_expected_worker_is_present():
    """FlinkSpoke.scala:145-156: a Create that arrives before the observed parallelism matches the expected one is
    parked (requestBuffer) and replayed when it does; records keep being buffered meanwhile."""
    from omldm_b200.engine import Engine
    from omldm_b200.stream import ring as R
    from omldm_b200.utils.config import JobParameters
    jp = JobParameters(parallelism=1, testSetSize=16)
    jp.extra["expectedParallelism"] = "2"
    e = Engine(jp, device="cpu")
    e.handle_request({"id": 1, "request": "Create", "learner": {"name": "PA"}, "trainingConfiguration": {"protocol": "Synchronous"}})
    assert not e.pipelines and len(e.pending_requests) == 1
    e.train_batch(R.generate_batch(R.KIND_BINARY, 5, 64, seed=1), 5)
    assert e._buffered_rows == 64
    e.observe_parallelism(2)
    assert list(e.pipelines) == [1] and not e.pending_requests
    e.train_batch(R.generate_batch(R.KIND_BINARY, 5, 64, seed=1, batch_index=1), 5)
    assert e._buffered_rows == 0 and e.count == 128


def test_default_test_mode_keeps_a_per_mini_batch_learning_curve():
    """FlinkHub.scala:101-113: in test mode (the default) the hub records (loss, #fitted) per mini-batch; the engine
    keeps those samples in a device ring, no host synchronisation per step."""
    from omldm_b200.engine import Engine
    from omldm_b200.stream import ring as R
    from omldm_b200.utils.config import JobParameters
    d = 8
    e = Engine(JobParameters(parallelism=1, testSetSize=16), device="cpu")          # test=True is the default
    e.handle_request({"id": 1, "request": "Create", "learner": {"name": "ORR", "hyperParameters": {"solveEvery": 1}},
                      "trainingConfiguration": {"protocol": "Synchronous"}})
    for i in range(12):
        e.train_batch(R.generate_batch(R.KIND_REGRESSION, d, 50, seed=4, batch_index=i), d)
    js = e.terminate()
    st = js.statistics[0]
    assert len(st.learningCurve) == len(st.lcx) >= 11 and st.lcx == sorted(st.lcx) and st.lcx[-1] == st.fitted
    assert st.learningCurve[-1] < st.learningCurve[1]            # the loss of a learnable stream goes down
    e2 = Engine(JobParameters(parallelism=1, testSetSize=16, test=False), device="cpu")
    e2.handle_request({"id": 1, "request": "Create", "learner": {"name": "PA"}, "trainingConfiguration": {"protocol": "Synchronous"}})
    e2.train_batch(R.generate_batch(R.KIND_BINARY, d, 50, seed=4), d)
    assert e2.pipelines[1].learner.curve_ring is None
