"""Regenerates tests/golden/*.json: JSON Request -> Response golden files (SURVEY §4 "API tests").

Each scenario replays a fixed request / record script through the fp64 host engine (deterministic: data from a
seeded generator, no threads) and stores every line the job wrote to the `responses` topic.  The test
(tests/test_golden.py) replays the script and compares keys, bucket ranges, bucket order, which fragment carries
the scalars and dataFitted exactly, and floating-point payloads to 1e-9.

    python scripts/make_golden.py          # rewrite the files after an intended format change
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))

from omldm_b200.runtime.job import HostJob                      # noqa: E402
from omldm_b200.utils.config import JobParameters               # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "..", "tests", "golden")

SCENARIOS = {
    # one fragment: everything fits one bucket
    "pa_small": dict(d=6, n=400, P=2, kind="cls",
                     create={"id": 3, "request": "Create", "learner": {"name": "PA", "hyperParameters": {"C": 0.1}},
                             "preProcessors": [{"name": "StandardScaler"}],
                             "trainingConfiguration": {"protocol": "Synchronous", "miniBatchSize": 8}}),
    # ORR with d = 120: A has 121 * 121 = 14 641 values -> keys a[0-9999] and a[10000-14640]; <= 2 buckets are sent whole
    "orr_two_buckets": dict(d=120, n=300, P=1, kind="reg",
                            create={"id": 5, "request": "Create",
                                    "learner": {"name": "ORR", "hyperParameters": {"lambda": 1.0}},
                                    "trainingConfiguration": {"protocol": "CentralizedTraining"}}),
    # ORR with d = 180: 181 * 181 = 32 761 values -> 4 parameter buckets, one fragment each; the last carries the scalars
    "orr_four_buckets": dict(d=180, n=260, P=1, kind="reg",
                             create={"id": 9, "request": "Create",
                                     "learner": {"name": "ORR", "hyperParameters": {"lambda": 0.5}},
                                     "trainingConfiguration": {"protocol": "CentralizedTraining"}}),
    # K-means is answered by worker 0 only (PipelineMap.scala:37-42)
    "kmeans": dict(d=4, n=500, P=2, kind="clu",
                   create={"id": 2, "request": "Create", "learner": {"name": "K-means", "hyperParameters": {"k": 3}},
                           "trainingConfiguration": {"protocol": "Asynchronous"}}),
}


def run_scenario(name: str):
    sc = SCENARIOS[name]
    rng = np.random.default_rng(abs(hash(name)) % (2 ** 31) if False else sum(map(ord, name)))
    d = sc["d"]
    w = rng.standard_normal(d) / np.sqrt(d)
    centres = rng.standard_normal((3, d)) * 4
    job = HostJob(JobParameters(parallelism=sc["P"], timeout=5, distributedKMeans=False))
    job.submit_request(sc["create"])
    for i in range(sc["n"]):
        x = rng.standard_normal(d)
        if sc["kind"] == "cls":
            y = float(np.sign(x @ w) or 1.0)
        elif sc["kind"] == "reg":
            y = float(x @ w + 0.25)
        else:
            c = int(rng.integers(0, 3))
            x = centres[c] + 0.3 * x
            y = float(c)
        job.submit_data({"numericalFeatures": [float(np.round(v, 6)) for v in x], "target": y, "operation": "training"})
    before = len(job.responses)
    job.submit_request({"id": sc["create"]["id"], "request": "Query", "requestId": 77})
    return [json.loads(r) for r in job.responses[before:]]


def digest(o):
    """Long numeric arrays are stored as a digest (length, sums, ends) so the golden files stay small."""
    if isinstance(o, dict):
        return {k: digest(v) for k, v in o.items()}
    if isinstance(o, list):
        if len(o) > 64 and all(isinstance(v, (int, float)) for v in o):
            a = np.asarray(o, dtype=np.float64)
            return {"__len__": len(o), "__sum__": float(a.sum()), "__abs__": float(np.abs(a).sum()),
                    "__head__": o[:3], "__tail__": o[-3:]}
        return [digest(v) for v in o]
    return o


def main() -> None:
    os.makedirs(GOLDEN, exist_ok=True)
    for name in SCENARIOS:
        frags = run_scenario(name)
        with open(os.path.join(GOLDEN, name + ".json"), "w") as f:
            json.dump({"scenario": name, "request": SCENARIOS[name]["create"], "responses": digest(frags)}, f, indent=1)
        print(name, len(frags), "fragment(s)")


if __name__ == "__main__":
    main()
