"""Topic-in / topic-out service (the reference's Kafka contract) on both engines, CPU."""
import json
import time

import numpy as np
import pytest

from omldm_b200.service import OnlineLearningService
from omldm_b200.utils.config import JobParameters


def _feed(svc, n, d=8, seed=0):
    rng = np.random.default_rng(seed)
    w = np.arange(1, d + 1) / d
    for _ in range(n):
        x = rng.standard_normal(d)
        svc.topic("trainingData").put({"numericalFeatures": x.tolist(), "target": float(np.sign(x @ w)),
                                       "operation": "training"})
    svc.topic("trainingData").put("EOS")


@pytest.mark.parametrize("engine,P", [("host", 2), ("gpu", 1)])
def test_service_topics_roundtrip_and_quiescence_termination(engine, P):
    params = JobParameters(parallelism=P, timeout=300, jobName="svc", testSetSize=32)
    svc = OnlineLearningService(params, engine=engine, device="cpu", batch=256).start()
    try:
        svc.topic("requests").put({"id": 1, "request": "Create", "learner": {"name": "PA", "hyperParameters": {"C": 0.1}},
                                   "trainingConfiguration": {"protocol": "Synchronous", "miniBatchSize": 16}})
        _feed(svc, 1500)
        deadline = time.time() + 20
        while len(svc.topic("trainingData")) and time.time() < deadline:
            time.sleep(0.01)
        svc.topic("forecastingData").put({"numericalFeatures": [0.1] * 8, "operation": "forecasting"})
        svc.topic("requests").put({"id": 1, "request": "Query", "requestId": 42})
        resp = None
        while time.time() < deadline and resp is None:
            resp = svc.topic("responses").get(timeout=0.2)
        assert resp is not None and svc.error is None, svc.error
        r = json.loads(resp)
        assert r["responseId"] == 42 and r["dataFitted"] > 1000 and r["score"] > 0.85
        pred = None
        while time.time() < deadline and pred is None:
            pred = svc.topic("predictions").get(timeout=0.2)
        assert pred is not None and json.loads(pred)["mlpId"] == 1
        # quiet input for `timeout` ms -> termination probe -> JobStatistics on `performance`
        assert svc.wait_terminated(20)
        perf = svc.topic("performance").get(timeout=1.0)
        assert perf is not None and json.loads(perf)["jobName"] == "svc"
        assert json.loads(perf)["statistics"][0]["fitted"] > 1000
    finally:
        svc.stop()


# ---- native JSON record scanner (csrc/host/json_records.cpp) --------------------------------------------------
def _parse(text, d, max_rows=1000, nt=1):
    from omldm_b200.stream.ring import parse_json_records
    return parse_json_records(text.encode() if isinstance(text, str) else text, d, max_rows, n_threads=nt)


def test_json_scanner_key_order_unknown_keys_and_escapes():
    lines = [
        '{"numericalFeatures":[1.5,-2,3e2],"target":4.25,"operation":"training"}',
        '{"operation":"forecasting","discreteFeatures":[7],"numericalFeatures":[0.125, 1E-3]}',          # discrete first
        '{"meta":{"a":[1,{"b":"}]"}],"s":"x\\\\\\"y"},"numericalFeatures":[1,2],"discreteFeatures":[3],"target":null}',
        '{"categoricalFeatures":["a,b]","numericalFeatures"],"numericalFeatures":[9,8,7],"target":-1}',   # key name inside a string value
        'EOS', '', '   ', '{"numericalFeatures":[1,2]}',                                                  # wrong width: dropped
        '{"numericalFeatures":[1,2,x]}', '{"numericalFeatures":[1,2,3],"operation":"delete"}',            # malformed / unknown op
        '{"numericalFeatures":[12345678901234567890123,1e-320,0.000001],"target":1e400}',
    ]
    rows, nf, used = _parse("\n".join(lines), 3)
    import numpy as np
    got = rows.numpy()
    assert rows.shape[0] == 5 and nf == 1 and used == len("\n".join(lines))
    np.testing.assert_array_equal(got[0, :5], np.float32([1.5, -2, 300, 1, 4.25]))
    np.testing.assert_array_equal(got[1, :5], np.float32([0.125, 1e-3, 7, 2, 0]))       # numerical, then discrete; flag 2
    np.testing.assert_array_equal(got[2, :5], np.float32([1, 2, 3, 1, 0]))
    np.testing.assert_array_equal(got[3, :5], np.float32([9, 8, 7, 1, -1]))
    assert got[4, 0] == np.float32(1.2345678901234568e22) and got[4, 1] == 0.0 and got[4, 2] == np.float32(1e-6)
    assert np.isinf(got[4, 4])


def test_json_scanner_threads_agree_and_numbers_round_like_numpy():
    import json
    import numpy as np
    import torch
    rng = np.random.default_rng(0)
    d, n = 12, 6000
    X = rng.standard_normal((n, d)) * 10.0 ** rng.integers(-8, 9, size=(n, d))
    y = rng.standard_normal(n)
    lines = []
    for i in range(n):
        if i % 97 == 0:
            lines.append("garbage line")
        lines.append(json.dumps({"numericalFeatures": X[i].tolist(), "target": float(y[i]),
                                 "operation": "forecasting" if i % 5 == 0 else "training"}))
    text = ("\n".join(lines) + "\n").encode()
    assert len(text) > (1 << 16)
    r1, f1, u1 = _parse(text, d, len(lines), 1)
    r4, f4, u4 = _parse(text, d, len(lines), 4)
    assert r1.shape[0] == n and f1 == f4 == len(range(0, n, 5)) and u1 == u4 == len(text)
    assert torch.equal(r1, r4)
    got, ref = r1[:, :d].numpy(), X.astype(np.float32)
    ulp = np.abs(got.view(np.int32).astype(np.int64) - ref.view(np.int32).astype(np.int64))
    assert ulp.max() <= 1 and (ulp == 0).mean() > 0.9999
    np.testing.assert_array_equal(r1[:, d + 1].numpy(), y.astype(np.float32))
    # max_rows: the single-thread scanner stops after max_rows ROWS, the threaded one after max_rows LINES;
    # both say where to resume and resuming yields the rest
    for nt in (1, 4):
        a, _, used = _parse(text, d, 1000, nt)
        b, _, _ = _parse(text[used:], d, len(lines), nt)
        assert a.shape[0] <= 1000 and a.shape[0] + b.shape[0] == n
        assert torch.equal(torch.cat([a, b]), r1)
