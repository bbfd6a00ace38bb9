"""Multi-rank streaming service (omldm_b200/dservice.py) on 2 ranks over gloo, CPU: requests arrive on rank 0 in the
middle of the stream and reach every worker at a step boundary (Create, Query, Delete, Create again), records that
arrive before the first Create are buffered and trained on afterwards, forecasts are answered from the local replica
between training batches, and the job ends by quiescence with JobStatistics on rank 0."""
import json
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _records(n, d, seed):
    rng = np.random.default_rng(seed)
    w = np.arange(1, d + 1) / d
    out = []
    for _ in range(n):
        x = rng.standard_normal(d)
        out.append(json.dumps({"numericalFeatures": x.tolist(), "target": float(np.sign(x @ w)), "operation": "training"}))
    return out


def _worker(rank, world, port, q, gpu=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    if gpu:       # one process per GPU, NCCL for the plumbing, the fused peer-memory protocol kernels for the steps
        import torch
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    from omldm_b200.dservice import DistributedService
    from omldm_b200.utils.config import JobParameters
    d = 8
    svc = DistributedService(JobParameters(parallelism=world, timeout=250, jobName="dsvc", testSetSize=16), rank, world,
                             f"cuda:{rank}" if gpu else "cpu", batch=128, impl="fused" if gpu else "nccl")
    log = {}
    # ---- records before any Create: buffered by the engine, nothing is trained
    svc.topic("trainingData").put_many(_records(300, d, 10 + rank))
    for _ in range(4):
        svc.step()
    log["buffered"] = svc.engine._buffered_rows
    log["pipelines_before"] = len(svc.engine.pipelines)
    # ---- Create in mid-stream (rank 0 only puts it; every rank gets it at the next round)
    if rank == 0:
        svc.topic("requests").put({"id": 1, "request": "Create", "learner": {"name": "PA", "hyperParameters": {"C": 0.1}},
                                   "trainingConfiguration": {"protocol": "Synchronous"}})
    svc.topic("trainingData").put_many(_records(500 + 60 * rank, d, 20 + rank))       # unequal partitions
    while True:
        svc.step()
        (left,) = svc._allmax([len(svc.topic("trainingData"))])
        if left == 0:
            break
    log["buffer_after_create"] = svc.engine._buffered_rows
    # ---- a forecast on rank 1 only, and a Query from rank 0, while the stream is still open
    if rank == 1:
        svc.topic("forecastingData").put({"numericalFeatures": [0.3] * d, "operation": "forecasting"})
    if rank == 0:
        svc.topic("requests").put({"id": 1, "request": "Query", "requestId": 7})
    svc.step()
    svc.step()
    log["responses"] = svc.topic("responses").drain()
    log["predictions"] = svc.topic("predictions").drain()
    log["fitted_local"] = svc.engine.pipelines[1].learner.counters()["fitted"]
    # ---- Delete, then a second pipeline on the same stream
    if rank == 0:
        svc.topic("requests").put({"id": 1, "request": "Delete"})
        svc.topic("requests").put({"id": 2, "request": "Create", "learner": {"name": "RegressorPA", "hyperParameters": {"C": 0.1}},
                                   "trainingConfiguration": {"protocol": "Asynchronous"}})
    svc.topic("trainingData").put_many(_records(256, d, 30 + rank))
    for _ in range(4):
        svc.step()
    log["pipelines_after"] = sorted(svc.engine.pipelines)
    log["steps"] = svc.steps
    # ---- quiescence: no input for `timeout` ms -> termination probe on every rank, JobStatistics on rank 0
    svc.run(max_rounds=2000, poll_ms=5.0)
    log["terminated"] = svc.terminated
    log["performance"] = svc.topic("performance").drain()
    svc.close()
    q.put((rank, log))
    dist.destroy_process_group()


def test_requests_in_mid_stream_on_two_ranks():
    _run_two_ranks(gpu=False)


@pytest.mark.gpu
def test_requests_in_mid_stream_on_two_gpus():
    """The same scenario with one process per GPU: NCCL bootstrap, fused protocol kernels over peer memory."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    _run_two_ranks(gpu=True)


def _run_two_ranks(gpu):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q, gpu)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    l0, l1 = res[0], res[1]
    # records that came before the Create waited in the buffer on both ranks, and were trained on after it
    assert l0["pipelines_before"] == l1["pipelines_before"] == 0
    assert l0["buffered"] >= 300 and l1["buffered"] >= 300
    assert l0["buffer_after_create"] == l1["buffer_after_create"] == 0
    assert l0["steps"] == l1["steps"]                                   # same number of equal-shape steps on both ranks
    # the Query was answered once, by rank 0, with the SUM of both workers' counts (buffered records included)
    assert len(l0["responses"]) == 1 and l1["responses"] == []
    r = json.loads(l0["responses"][0])
    assert r["responseId"] == 7 and r["protocol"] == "Synchronous"
    assert r["dataFitted"] == l0["fitted_local"] + l1["fitted_local"]
    assert r["dataFitted"] > 2 * 700 and r["score"] > 0.8
    # the forecast was answered by the rank that received it, from its local replica
    assert l0["predictions"] == [] and len(l1["predictions"]) == 1 and json.loads(l1["predictions"][0])["mlpId"] == 1
    # Delete + second Create took effect on both ranks
    assert l0["pipelines_after"] == l1["pipelines_after"] == [2]
    # quiescence ended the job on both ranks; rank 0 published the statistics
    assert l0["terminated"] and l1["terminated"]
    assert len(l0["performance"]) == 1 and l1["performance"] == []
    js = json.loads(l0["performance"][0])
    assert js["jobName"] == "dsvc" and js["parallelism"] == 2 and js["statistics"][0]["pipeline"] == 2
