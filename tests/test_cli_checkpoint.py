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


def _two_worker_checkpoint(tmp_path, learner, pre=None, d=7, kind=None):
    """Two independent workers (the two subtasks of a parallelism-2 job between synchronisations)
    write rank000.pt / rank001.pt plus a manifest that says world = 2."""
    kind = R.KIND_REGRESSION if kind is None else kind
    engines, data = [], []
    for r in range(2):
        e = Engine(JobParameters(parallelism=1, testSetSize=16), device="cpu")
        e.handle_request({"id": 1, "request": "Create", "learner": learner, "preProcessors": pre or [],
                          "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
        bs = [R.generate_batch(kind, d, 64, seed=20 + r, batch_index=i) for i in range(3)]
        for b in bs:
            e.train_batch(b.clone(), d)
        e.rank = r
        save_checkpoint(e, str(tmp_path / "ck"), step=3)
        engines.append(e)
        data.append(bs)
    man = json.load(open(tmp_path / "ck" / "manifest.json"))
    man["world"] = 2
    json.dump(man, open(tmp_path / "ck" / "manifest.json", "w"))
    return engines, data


def test_restore_with_fewer_ranks_merges_worker_states(tmp_path):
    """Rescaling restore (FlinkSpoke.scala:289-305): one new worker folds both old ones."""
    d = 7
    engines, _ = _two_worker_checkpoint(tmp_path, {"name": "ORR", "hyperParameters": {"lambda": 1.0}},
                                        [{"name": "StandardScaler"}], d)
    m = Engine(JobParameters(parallelism=1, testSetSize=16), device="cpu")
    assert load_checkpoint(m, str(tmp_path / "ck")) == 3
    la, lb, lm = engines[0].pipelines[1].learner, engines[1].pipelines[1].learner, m.pipelines[1].learner
    # additive statistics: the merged Gram delta is the sum of both workers' deltas
    assert torch.allclose(lm.buf, la.buf + lb.buf, rtol=1e-5, atol=1e-5)
    ca, cb, cm = la.counters(), lb.counters(), lm.counters()
    assert cm["fitted"] == ca["fitted"] + cb["fitted"]
    assert abs(cm["loss_sum"] - ca["loss_sum"] - cb["loss_sum"]) < 1e-6 * max(1.0, abs(cm["loss_sum"]))
    # scaler moments merge exactly (Chan): count adds, mean is the weighted mean
    sa, sb, sm = (e.pipelines[1].preprocessors[0].state for e in (*engines, m))
    assert float(sm[0]) == float(sa[0]) + float(sb[0])
    wmean = (sa[0] * sa[1:1 + d] + sb[0] * sb[1:1 + d]) / (sa[0] + sb[0])
    assert torch.allclose(sm[1:1 + d], wmean, rtol=1e-10, atol=1e-12)
    # holdout rings interleave, nothing beyond the filled rows counts as valid
    assert m.holdout.shape == engines[0].holdout.shape
    n_valid = m._holdout_rows().shape[0]
    assert 0 < n_valid <= 16 and bool((m._holdout_rows()[:, d] == 1.0).all())
    # and the merged worker keeps training
    m.train_batch(R.generate_batch(R.KIND_REGRESSION, d, 64, seed=99, batch_index=0), d)
    assert m.pipelines[1].learner.counters()["fitted"] > cm["fitted"]


def test_restore_averaged_model_and_more_ranks(tmp_path):
    d = 7
    engines, _ = _two_worker_checkpoint(tmp_path, {"name": "PA", "hyperParameters": {"C": 0.1}}, None, d,
                                        kind=R.KIND_BINARY)
    m = Engine(JobParameters(parallelism=1, testSetSize=16), device="cpu")
    load_checkpoint(m, str(tmp_path / "ck"))
    wa, wb, wm = (e.pipelines[1].learner.w for e in (*engines, m))
    assert torch.allclose(wm, 0.5 * (wa + wb), atol=1e-6)
    # a rank that did not exist before adopts the model of rank (r mod old_world) with empty history
    g = Engine(JobParameters(parallelism=1, testSetSize=16), device="cpu")
    g.rank, g.world = 3, 4
    load_checkpoint(g, str(tmp_path / "ck"))
    assert torch.allclose(g.pipelines[1].learner.w, wb, atol=1e-7)
    assert g.pipelines[1].learner.counters()["fitted"] == 0 and g.count == 0
    assert g._holdout_rows().shape[0] == 0


def test_cli_resume_starts_at_the_stream_position_and_requests_apply_in_mid_stream(tmp_path):
    """A run cut in two by a checkpoint + `--resume true` trains every record exactly once (the topic restarts at the
    checkpointed step), and a request with "atBatch" takes effect at that step boundary, not before the stream."""
    req = tmp_path / "req.jsonl"
    with open(req, "w") as f:
        f.write(json.dumps({"id": 1, "request": "Create", "learner": {"name": "PA", "hyperParameters": {"C": 0.1}},
                            "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": 12}}) + "\n")
        f.write(json.dumps({"id": 2, "request": "Create", "atBatch": 6, "learner": {"name": "RegressorPA"},
                            "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": 12}}) + "\n")
        f.write(json.dumps({"id": 1, "request": "Query", "requestId": 5}) + "\n")
        f.write(json.dumps({"id": 2, "request": "Query", "requestId": 6}) + "\n")
    common = ["--engine", "gpu", "--requests", str(req), "--trainingData", "synthetic:binary:12", "--batch", "100",
              "--testSetSize", "20", "--test", "false"]
    # one uninterrupted run of 10 batches
    out_a = tmp_path / "a"
    assert cli.main(common + ["--batches", "10", "--outDir", str(out_a)]) == 0
    ra = {json.loads(l)["mlpId"]: json.loads(l) for l in open(out_a / "responses.jsonl")}
    assert ra[1]["dataFitted"] == 10 * 100 - 20
    assert ra[2]["dataFitted"] == 4 * 100           # created at the boundary before step 6: sees batches 6..9 only
    # the same stream cut in two: 4 batches with a checkpoint at the end, then resume
    ck = tmp_path / "ck"
    out_b = tmp_path / "b"
    assert cli.main(common + ["--batches", "4", "--outDir", str(out_b), "--checkpointing", "true", "--checkInterval", "0",
                              "--stateBackend", str(ck)]) == 0
    assert json.load(open(ck / "manifest.json"))["step"] == 4
    out_c = tmp_path / "c"
    assert cli.main(common + ["--batches", "10", "--outDir", str(out_c), "--checkpointing", "true", "--checkInterval",
                              "100000000", "--stateBackend", str(ck), "--resume", "true"]) == 0
    rc = {json.loads(l)["mlpId"]: json.loads(l) for l in open(out_c / "responses.jsonl")}
    assert rc[1]["dataFitted"] == ra[1]["dataFitted"]          # nothing trained twice, nothing skipped
    assert rc[2]["dataFitted"] == ra[2]["dataFitted"]
    wa = ra[1]["learner"]["parameters"]["a"]
    wc = rc[1]["learner"]["parameters"]["a"]
    assert np.allclose(np.asarray(wa, dtype=float), np.asarray(wc, dtype=float), rtol=1e-5, atol=1e-6)
