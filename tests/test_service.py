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
