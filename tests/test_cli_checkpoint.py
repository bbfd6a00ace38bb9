"""CLI (both engines, CPU) and checkpoint / resume."""
import json
import os

import numpy as np
import torch

from omldm_b200 import cli
from omldm_b200.engine import Engine
from omldm_b200.stream import ring as R
from omldm_b200.utils.checkpoint import PeriodicCheckpointer, load_checkpoint, save_checkpoint
from omldm_b200.utils.config import JobParameters


def _requests(path):
    with open(path, "w") as f:
        f.write(json.dumps({"id": 1, "request": "Create", "learner": {"name": "PA", "hyperParameters": {"C": 0.1}},
                            "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": 6}}) + "\n")
        f.write(json.dumps({"id": 1, "request": "Query", "requestId": 11}) + "\n")


def _records(path, n, op="training"):
    rng = np.random.default_rng(0)
    w = rng.standard_normal(6)
    with open(path, "w") as f:
        for _ in range(n):
            x = rng.standard_normal(6)
            rec = {"numericalFeatures": x.tolist(), "operation": op}
            if op == "training":
                rec["target"] = float(np.sign(x @ w))
            f.write(json.dumps(rec) + "\n")
        f.write("EOS\n")


def test_flag_parsing_matches_reference_names():
    p = JobParameters.from_args(["--parallelism", "8", "--jobName", "j", "--test", "false", "--maxMsgParams", "500",
                                 "--timeout", "100", "--testSetSize", "64", "--trainingDataTopic", "td",
                                 "--requestsAddr", "host:9092", "--checkpointing", "true", "--checkInterval", "10"])
    assert (p.parallelism, p.jobName, p.test, p.maxMsgParams, p.timeout, p.testSetSize) == (8, "j", False, 500, 100, 64)
    assert p.topics["trainingData"] == "td" and p.addrs["requests"] == "host:9092" and p.checkpointing
    d = JobParameters()
    assert (d.parallelism, d.maxMsgParams, d.test, d.timeout, d.testSetSize, d.jobName) == \
        (16, 2000, True, 30000, 256, "OML_job_1")


def test_cli_host_and_gpu_engines_on_cpu(tmp_path):
    req, tr, fc = tmp_path / "req.jsonl", tmp_path / "train.jsonl", tmp_path / "fc.jsonl"
    _requests(req); _records(tr, 600); _records(fc, 5, "forecasting")
    out1 = tmp_path / "out_host"
    assert cli.main(["--engine", "host", "--parallelism", "2", "--requests", str(req), "--trainingData", str(tr),
                     "--forecastingData", str(fc), "--outDir", str(out1), "--timeout", "5"]) == 0
    r = json.loads(open(out1 / "responses.jsonl").readline())
    assert r["responseId"] == 11 and r["score"] > 0.8
    assert len(open(out1 / "predictions.jsonl").readlines()) == 5
    assert json.loads(open(out1 / "performance.jsonl").readline())["parallelism"] == 2
    out2 = tmp_path / "out_gpu"
    assert cli.main(["--engine", "gpu", "--requests", str(req), "--trainingData", str(tr), "--forecastingData", str(fc),
                     "--batch", "128", "--outDir", str(out2)]) == 0
    r2 = json.loads(open(out2 / "responses.jsonl").readline())
    assert r2["responseId"] == 11 and r2["dataFitted"] > 0 and r2["score"] > 0.8
    assert len(open(out2 / "predictions.jsonl").readlines()) == 5


def test_checkpoint_resume_continues_identically(tmp_path):
    d = 9
    def mk():
        e = Engine(JobParameters(parallelism=1, testSetSize=32), device="cpu")
        e.handle_request({"id": 1, "request": "Create", "learner": {"name": "RegressorPA", "hyperParameters": {"C": 0.5}},
                          "preProcessors": [{"name": "StandardScaler"}],
                          "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
        return e
    batches = [R.generate_batch(R.KIND_REGRESSION, d, 128, seed=8, batch_index=i) for i in range(8)]
    a = mk()
    for b in batches[:4]:
        a.train_batch(b.clone(), d)
    save_checkpoint(a, str(tmp_path / "ck"), step=4)
    assert json.load(open(tmp_path / "ck" / "manifest.json"))["step"] == 4
    for b in batches[4:]:
        a.train_batch(b.clone(), d)
    b_eng = Engine(JobParameters(parallelism=1, testSetSize=32), device="cpu")
    assert load_checkpoint(b_eng, str(tmp_path / "ck")) == 4
    for b in batches[4:]:
        b_eng.train_batch(b.clone(), d)
    wa = a.pipelines[1].learner.w
    wb = b_eng.pipelines[1].learner.w
    assert torch.allclose(wa, wb, atol=1e-6) and a.count == b_eng.count
    pc = PeriodicCheckpointer(a, str(tmp_path / "ck2"), interval_ms=0)
    assert pc.maybe_save(9) and os.path.exists(tmp_path / "ck2" / "rank000.pt")
