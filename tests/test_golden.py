"""JSON Request -> Response golden files (SURVEY §4 "API tests"): bucket keys `name[start-end]`, bucket size
10 000, fragment order, which fragment carries the scalars (FlinkNetwork.scala:48-240), and the payload values.
Regenerate with `python scripts/make_golden.py` after an intended format change."""
import json
import math
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "scripts"))
import make_golden  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _same(a, b, path=""):
    if isinstance(a, dict):
        assert isinstance(b, dict) and list(a.keys()) == list(b.keys()), f"{path}: keys {list(a)} != {list(b)}"
        for k in a:
            _same(a[k], b[k], f"{path}/{k}")
    elif isinstance(a, list):
        assert isinstance(b, list) and len(a) == len(b), f"{path}: length"
        for i, (x, y) in enumerate(zip(a, b)):
            _same(x, y, f"{path}[{i}]")
    elif isinstance(a, float) or isinstance(b, float):
        assert a is not None and b is not None, path
        assert math.isclose(a, b, rel_tol=1e-9, abs_tol=1e-9), f"{path}: {a} != {b}"
    else:
        assert a == b, f"{path}: {a!r} != {b!r}"


@pytest.mark.parametrize("name", sorted(make_golden.SCENARIOS))
def test_responses_match_golden(name):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        gold = json.load(f)
    live = make_golden.digest(make_golden.run_scenario(name))
    assert len(live) == len(gold["responses"])
    _same(gold["responses"], json.loads(json.dumps(live)))


def test_golden_bucket_contract():
    """What the files pin down, stated explicitly."""
    g = json.load(open(os.path.join(GOLDEN, "orr_four_buckets.json")))["responses"]
    assert [f["bucketIndex"] for f in g] == [0, 1, 2, 3]
    keys = [sorted(f["learner"]["parameters"]) for f in g]
    assert keys[0] == ["a[0-9999]", "b", "weights"] and keys[1] == ["a[10000-19999]"]
    assert keys[2] == ["a[20000-29999]"] and keys[3] == ["a[30000-32760]"]
    assert all(f["learner"]["parameters"][k]["__len__"] == 10000 for f, k in zip(g[:3], ("a[0-9999]", "a[10000-19999]", "a[20000-29999]")))
    for f in g[:3]:                                   # only the last parameter bucket carries the scalars
        assert f["dataFitted"] is None and f["loss"] is None and f["score"] is None and f["protocol"] is None
    assert g[3]["dataFitted"] == 208 and g[3]["protocol"] == "CentralizedTraining" and g[3]["score"] is not None
    two = json.load(open(os.path.join(GOLDEN, "orr_two_buckets.json")))["responses"]
    assert len(two) == 1 and two[0]["bucketIndex"] is None          # <= 2 buckets: sent whole (:187-191)
    assert two[0]["learner"]["parameters"]["a"]["__len__"] == 121 * 121
    km = json.load(open(os.path.join(GOLDEN, "kmeans.json")))["responses"]
    assert len(km) == 1 and km[0]["protocol"] == "SingleLearner"     # forced for K-means (FlinkSpoke.scala:207-209)


def test_gpu_engine_answers_in_the_same_layout():
    """The GPU engine (here on its CPU path) fragments an ORR answer exactly like the host engine's golden file."""
    from omldm_b200.engine import Engine
    from omldm_b200.stream import ring as R
    from omldm_b200.utils.config import JobParameters
    g = json.load(open(os.path.join(GOLDEN, "orr_four_buckets.json")))
    d = 180
    eng = Engine(JobParameters(parallelism=1), device="cpu")       # default holdout capacity (256), like the golden run
    eng.handle_request(dict(g["request"], trainingConfiguration=dict(g["request"]["trainingConfiguration"], nFeatures=d)))
    eng.train_batch(R.generate_batch(R.KIND_REGRESSION, d, 260, seed=1), d)
    frags = [json.loads(s) for s in eng.handle_request({"id": 9, "request": "Query", "requestId": 77})]
    assert [f["bucketIndex"] for f in frags] == [f["bucketIndex"] for f in g["responses"]]
    for mine, gold in zip(frags, g["responses"]):
        assert list(mine.keys()) == list(gold.keys())
        assert sorted(mine["learner"]["parameters"]) == sorted(gold["learner"]["parameters"])
        assert (mine["dataFitted"] is None) == (gold["dataFitted"] is None)
    assert frags[-1]["dataFitted"] == 208 and frags[-1]["protocol"] == "CentralizedTraining"
