"""Developer probe: where does the ORR step time go (graph replay vs solve vs eager)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from omldm_b200.engine import Engine
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters
from omldm_b200.ops import functional as F

def ev_time(fn, n=20, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3, (time.perf_counter() - t0) / n * 1e6

dev = torch.device("cuda", 0)
d, B = 1000, 16384
for graphs in ("true", "false"):
    p = JobParameters(parallelism=1); p.extra["graphs"] = graphs
    eng = Engine(p, 0, 1, dev)
    eng.handle_request({"id": 1, "request": "Create", "learner": {"name": "ORR", "hyperParameters": {"solveEvery": 1000000}},
                        "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
    pool = [R.generate_batch(R.KIND_REGRESSION, d, B, seed=1, batch_index=i).to(dev) for i in range(5)]
    i = [0]
    def step():
        eng.train_batch(pool[i[0] % 5], d); i[0] += 1
    g, w = ev_time(step, 40, 12)
    print(f"graphs={graphs}: step gpu {g:.1f} us, wall {w:.1f} us", flush=True)
    lrn = eng.pipelines[1].learner
    g, w = ev_time(lrn.solve, 5, 2)
    print(f"  solve gpu {g:.1f} us wall {w:.1f} us", flush=True)
    Z = pool[0]
    for name, fn in [("holdout", lambda: F.holdout_swap_dev(Z, eng.holdout, d, eng._count_dev)),
                     ("sq_residual", lambda: F.sq_residual(Z, d, lrn.wt, lrn.stats)),
                     ("gram", lambda: F.gemm_outer(Z, Z, lrn.dG, accumulate=True, ksplit=7, upper_only=True, m=d+2, n=d+2)),
                     ("fold", lambda: (lrn.e.add_(lrn.buf), lrn.buf.zero_()))]:
        g, w = ev_time(fn, 20, 3)
        print(f"  {name}: gpu {g:.1f} us wall {w:.1f} us", flush=True)
    eng.close()
