"""Multi-process (gloo, world_size 2) run of the engine's distributed host-side logic on CPU: the
torch.distributed protocol drivers (the NCCL-baseline code path), query merging, SingleLearner
forwarding and termination statistics."""
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


def _worker(rank, world, port, proto, learner, hyper, kind, q, tc_extra=None):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from omldm_b200.engine import Engine
    from omldm_b200.stream import ring as R
    from omldm_b200.utils.config import JobParameters
    d = 10
    eng = Engine(JobParameters(parallelism=world, testSetSize=16), rank, world, "cpu", impl="nccl")
    eng.handle_request({"id": 1, "request": "Create", "learner": {"name": learner, "hyperParameters": hyper},
                        "trainingConfiguration": dict({"protocol": proto, "nFeatures": d, "threshold": 0.01},
                                                      **(tc_extra or {}))})
    for i in range(12):
        eng.train_batch(R.generate_batch(kind, d, 64, seed=3, batch_index=2 * i + rank, n_aux=3), d)
    out = eng.handle_request({"id": 1, "request": "Query", "requestId": 4})
    js = eng.terminate()
    p = eng.pipelines[1]
    flat = (p.learner.e.clone() if getattr(p.driver, "delta", None) is not None else p.comm.buf.clone()) if p.built else None
    q.put((rank, out, None if js is None else js.to_dict(), None if flat is None else flat.numpy()))
    dist.destroy_process_group()


def _launch(proto, learner, hyper, kind, tc_extra=None):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, proto, learner, hyper, kind, q, tc_extra)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
    return res


@pytest.mark.parametrize("proto", ["Synchronous", "Asynchronous", "FGM"])
def test_two_workers_linear(proto):
    res = _launch(proto, "PA", {"C": 0.1}, 0)
    (r0, out0, js0, f0), (r1, out1, js1, f1) = res
    assert out1 == [] and len(out0) == 1                    # rank 0 answers for the job
    resp = json.loads(out0[0])
    assert resp["protocol"] == proto and resp["dataFitted"] == 2 * (12 * 64 - 16)   # SUM over workers
    assert 0.0 <= resp["score"] <= 1.0 and resp["score"] > 0.7
    if proto == "Synchronous":
        assert np.allclose(f0, f1)                          # replicas identical after every step
    assert js1 is None and js0["parallelism"] == 2 and js0["statistics"][0]["fitted"] == resp["dataFitted"]


def test_two_workers_orr_sum_merge_and_single_learner():
    res = _launch("Synchronous", "ORR", {"lambda": 1.0, "solveEvery": 1}, 1)
    resp = json.loads(res[0][1][0]) if len(res[0][1]) == 1 else None
    frags = [json.loads(o) for o in res[0][1]]
    assert frags and frags[-1]["dataFitted"] == 2 * (12 * 64 - 16) if len(frags) > 1 else resp["dataFitted"] > 0
    res = _launch("Synchronous", "HT", {"nClasses": 2, "gracePeriod": 50}, 0)
    (r0, out0, js0, _), (r1, out1, js1, _) = res
    resp = json.loads(out0[0])
    assert resp["protocol"] == "SingleLearner" and out1 == []
    assert resp["dataFitted"] == 2 * (12 * 64 - 16)         # both workers forwarded to the one learner


def test_scatter_fused_synchronisation_matches_plain_synchronous():
    """`syncFusion: "scatter"` (private delta + three rotating symmetric copies) on the torch.distributed path:
    same global statistics and the same answer as the plain Synchronous driver."""
    plain = _launch("Synchronous", "ORR", {"lambda": 1.0, "solveEvery": 1}, 1)
    fused = _launch("Synchronous", "ORR", {"lambda": 1.0, "solveEvery": 1}, 1, {"syncFusion": "scatter"})
    assert np.allclose(fused[0][3], fused[1][3])                       # both replicas hold the same global matrix
    fp, ff = [json.loads(o) for o in plain[0][1]], [json.loads(o) for o in fused[0][1]]
    assert len(fp) == len(ff) and fp[-1]["dataFitted"] == ff[-1]["dataFitted"] == 2 * (12 * 64 - 16)
    assert abs(fp[-1]["score"] - ff[-1]["score"]) < 1e-4 and abs(fp[-1]["cumulativeLoss"] - ff[-1]["cumulativeLoss"]) < 1e-3
