"""Host (fp64, message-driven) runtime: protocol invariants on a CPU-only box, request life cycle,
statistics / termination, checkpoint / restore (SURVEY §4 'protocol tests without GPUs')."""
import json

import numpy as np
import pytest

from omldm_b200.parallel.host_protocols import PROTOCOLS, MLNodeGenerator, shard_bounds
from omldm_b200.parallel.network import DataSet
from omldm_b200.runtime.job import HostJob
from omldm_b200.utils.config import JobParameters

RNG = np.random.default_rng(1)
W = RNG.standard_normal(12)


def rec(x, y, op="training"):
    d = {"numericalFeatures": [float(v) for v in x], "operation": op}
    if y is not None:
        d["target"] = float(y)
    return d


def make_job(proto, P=4, learner="PA", hyper=None, tc=None, **kw):
    job = HostJob(JobParameters(parallelism=P, timeout=5, **kw))
    cfg = dict({"protocol": proto, "miniBatchSize": 8, "HubParallelism": 2, "threshold": 0.05}, **(tc or {}))
    job.submit_request({"id": 1, "request": "Create", "learner": {"name": learner, "hyperParameters": hyper or {"C": 0.1}},
                        "trainingConfiguration": cfg})
    return job


def feed(job, n, kind="cls"):
    for _ in range(n):
        x = RNG.standard_normal(12)
        y = np.sign(x @ W) if kind == "cls" else x @ W + 0.5
        job.submit_data(rec(x, y))


@pytest.mark.parametrize("proto", ["Asynchronous", "Synchronous", "SSP", "EASGD", "GM", "FGM", "SingleLearner"])
def test_every_protocol_trains_and_answers(proto):
    job = make_job(proto)
    feed(job, 3000)
    job.submit_request({"id": 1, "request": "Query", "requestId": 7})
    r = json.loads(job.responses[-1])
    assert r["responseId"] == 7 and r["protocol"] == proto and r["score"] > 0.85
    js = job.run()
    assert js is not None and js.statistics[0].protocol == proto and js.statistics[0].fitted > 0
    assert job.terminated and json.loads(job.performance[-1])["parallelism"] == 4


def test_synchronous_is_exact_mean_of_replicas():
    job = make_job("Synchronous", tc={"HubParallelism": 3})
    feed(job, 4 * 8 * 5)   # 5 full rounds of 8 trained records per worker ... minus holdouts
    flats = [s.state[1].node.flat.copy() for s in job.spokes]
    es = [s.state[1].node.e.copy() for s in job.spokes]
    for e in es[1:]:
        assert np.allclose(e, es[0])                # every worker installed the same global model
    hub_global = np.concatenate([job.hubs[f"1_{h}"].state.get_node().global_shard for h in range(3)])
    assert np.allclose(hub_global, es[0])


def test_ssp_staleness_bound_is_respected():
    job = make_job("SSP", tc={"staleness": 1, "HubParallelism": 1})
    # feed worker 0 only: it may run at most `staleness` clocks ahead, then it blocks and buffers
    for _ in range(400):
        x = RNG.standard_normal(12)
        job.submit_data(rec(x, np.sign(x @ W)), worker=0)
    w0 = job.spokes[0].state[1]
    hub = job.hubs["1_0"].state.get_node()
    assert w0.node.blocked and len(w0.buffer) > 0
    assert hub.clocks[0] - hub.min_clock() == 2      # one clock beyond the bound: reply deferred
    for w in (1, 2, 3):                              # the stragglers catch up -> worker 0 is released
        for _ in range(40):
            x = RNG.standard_normal(12)
            job.submit_data(rec(x, np.sign(x @ W)), worker=w)
    # worker 0 was released, drained its backlog and ran ahead again - but never past the bound
    assert hub.clocks[0] > 2
    assert hub.clocks[0] - hub.min_clock() <= 2


def test_fgm_rounds_and_subrounds_happen_and_gm_rounds():
    job = make_job("FGM", tc={"threshold": 0.02, "HubParallelism": 1})
    feed(job, 4000)
    hub = job.hubs["1_0"].state.get_node()
    assert hub.rounds >= 1 and hub.subrounds >= 1
    job2 = make_job("GM", tc={"threshold": 0.02, "HubParallelism": 1})
    feed(job2, 2000)
    assert job2.hubs["1_0"].state.get_node().rounds >= 1


def test_parallelism_one_forces_centralized_and_ht_kmeans_single_learner():
    job = make_job("Synchronous", P=1)
    feed(job, 200)
    job.submit_request({"id": 1, "request": "Query", "requestId": 1})
    assert json.loads(job.responses[-1])["protocol"] == "CentralizedTraining"
    job = HostJob(JobParameters(parallelism=3, distributedKMeans=False))
    job.submit_request({"id": 2, "request": "Create", "learner": {"name": "K-means", "hyperParameters": {"k": 3}},
                        "trainingConfiguration": {"protocol": "Synchronous"}})
    for _ in range(300):
        job.submit_data(rec(RNG.standard_normal(4), None))
    job.submit_request({"id": 2, "request": "Query", "requestId": 2})
    r = json.loads(job.responses[-1])
    assert r["protocol"] == "SingleLearner" and r["dataFitted"] == 240


def test_unknown_protocol_falls_back_to_asynchronous():
    from omldm_b200.api.pojos import Request
    req = Request.from_dict({"id": 1, "request": "Create", "learner": {"name": "PA"},
                             "trainingConfiguration": {"protocol": "Gossip"}})
    g = MLNodeGenerator()
    assert g.generate_spoke_node(req).protocol == "Asynchronous"
    assert g.generate_hub_node(req).protocol == "Asynchronous"
    assert len(PROTOCOLS) == 8


def test_multi_pipeline_update_delete_and_forecast():
    job = make_job("Asynchronous")
    job.submit_request({"id": 2, "request": "Create", "learner": {"name": "RegressorPA"},
                        "preProcessors": [{"name": "PolynomialFeatures"}, {"name": "StandardScaler"}],
                        "trainingConfiguration": {"protocol": "Synchronous"}})
    feed(job, 800)
    job.submit_request({"id": 1, "request": "Update"})                               # accepted, no-op
    job.submit_data(rec(RNG.standard_normal(12), None, "forecasting"))
    assert len(job.predictions) == 2                                                 # every pipeline predicts
    assert {json.loads(p)["mlpId"] for p in job.predictions} == {1, 2}
    job.submit_request({"id": 1, "request": "Delete"})
    job.submit_data(rec(RNG.standard_normal(12), None, "forecasting"))
    assert len(job.predictions) == 3
    job.submit_request({"id": 1, "request": "Query", "requestId": 5})
    assert not any(json.loads(r)["mlpId"] == 1 and json.loads(r)["responseId"] == 5 for r in job.responses)


def test_records_are_buffered_until_a_pipeline_exists():
    job = HostJob(JobParameters(parallelism=2))
    for _ in range(50):
        x = RNG.standard_normal(12)
        job.submit_data(rec(x, np.sign(x @ W)))
    assert sum(len(s.record_buffer) for s in job.spokes) == 50
    job.submit_request({"id": 1, "request": "Create", "learner": {"name": "PA"},
                        "trainingConfiguration": {"protocol": "Asynchronous"}})
    x = RNG.standard_normal(12)
    job.submit_data(rec(x, 1.0)); job.submit_data(rec(x, 1.0))
    assert sum(len(s.record_buffer) for s in job.spokes) == 0


def test_checkpoint_restore_and_rescale():
    job = make_job("Synchronous", P=4)
    feed(job, 1200)
    snap = job.snapshot()
    job2 = HostJob(JobParameters(parallelism=2, timeout=5))                          # rescale 4 -> 2
    job2.restore(snap)
    assert 1 in job2.spokes[0].state and 1 in job2.spokes[1].state
    a = job.spokes[0].state[1].node.flat
    assert np.allclose(job2.spokes[0].state[1].node.flat, a)
    feed(job2, 200)
    job2.submit_request({"id": 1, "request": "Query", "requestId": 3})
    assert json.loads(job2.responses[-1])["score"] > 0.85


def test_dataset_bounded_fifo_and_shards():
    ds = DataSet(3)
    assert [ds.append(i) for i in range(5)] == [None, None, None, 0, 1]
    assert ds.pop() == 2 and len(ds) == 2
    other = DataSet(3); other.append(10); other.append(11)
    assert list(ds.merge([other]).data_buffer) == [10, 4, 11]
    assert shard_bounds(10, 3) == [(0, 4), (4, 8), (8, 10)]
