"""Timed CPU baseline: the fp64 message-driven host engine (stand-in for the reference's Flink
local mini-cluster) on the five BASELINE.json configs, scaled down to CPU-feasible sizes."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from omldm_b200.runtime.job import HostJob
from omldm_b200.stream.points import TrainingPoint
from omldm_b200.utils.config import JobParameters

rng = np.random.default_rng(0)
CFG = [
    ("cfg1 PA 28-d async P=4", 4, "PA", {"C": 0.01}, [], {"protocol": "Asynchronous", "miniBatchSize": 64}, 28, 40000, "cls"),
    ("cfg2 ORR 1000-d sync P=8 (d=200 on CPU)", 8, "ORR", {"lambda": 1.0, "solveEvery": 64}, [], {"protocol": "Synchronous", "miniBatchSize": 64}, 200, 4000, "reg"),
    ("cfg3 RegressorPA+poly+scaler FGM P=8", 8, "RegressorPA", {"C": 0.01}, [{"name": "PolynomialFeatures"}, {"name": "StandardScaler"}],
     {"protocol": "FGM", "miniBatchSize": 64, "threshold": 0.05}, 28, 8000, "reg"),
    ("cfg4 K-means k=256 128-d SSP P=8", 8, "K-means", {"k": 256}, [], {"protocol": "SSP", "miniBatchSize": 64}, 128, 8000, "clu"),
    ("cfg5 MultiClassPA 100k-d (d=2000 on CPU) sync P=8", 8, "MultiClassPA", {"nClasses": 10}, [], {"protocol": "Synchronous", "miniBatchSize": 64, "HubParallelism": 8}, 2000, 2000, "mc"),
]
for name, P, learner, hyper, pre, tc, d, n, kind in CFG:
    job = HostJob(JobParameters(parallelism=P, timeout=5, test=False, distributedKMeans=True))
    job.submit_request({"id": 1, "request": "Create", "learner": {"name": learner, "hyperParameters": hyper}, "preProcessors": pre, "trainingConfiguration": tc})
    w = rng.standard_normal(d)
    X = rng.standard_normal((n, d))
    y = np.sign(X @ w) if kind == "cls" else (X @ w if kind == "reg" else (rng.integers(0, 10, n).astype(float) if kind == "mc" else np.zeros(n)))
    pts = [TrainingPoint(X[i], target=float(y[i])) for i in range(n)]
    t0 = time.perf_counter()
    for p in pts:
        job.submit_data(p)
    dt = time.perf_counter() - t0
    print(json.dumps({"config": name, "records": n, "seconds": round(dt, 3), "records_per_s": round(n / dt, 1)}), flush=True)
