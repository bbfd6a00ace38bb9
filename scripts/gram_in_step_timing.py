"""Phase cycles of the Gram kernel INSIDE the real ORR step (beside the loss pass), against the kernel on its own.
Needs a tuning build: scripts/build_variant.sh timing csrc/gram_f16.cu -DOMLDM_GRAM_TIMING, then
OMLDM_KERNELS_LIB=omldm_b200/lib/libvar_timing.so python scripts/gram_in_step_timing.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch

from omldm_b200.engine import Engine
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters
from test_gram_f16_gpu import gram_phase_report



def timeline():
    """Where holdout / Gram / loss of the LAST step sit in time (tuning build with -DOMLDM_TIMELINE)."""
    import ctypes
    import numpy as np
    from omldm_b200.ops import _lib
    lib = _lib.kernels()
    if not hasattr(lib, "omldm_tl_loss"):
        return
    torch.cuda.synchronize()
    buf = np.zeros(160 * 16, dtype=np.int64)
    lib.omldm_gram_f16_stamps(buf.ctypes.data_as(ctypes.c_void_p))
    st = buf.reshape(160, 16)[:140]
    g0, g1 = int(st[:, 8][st[:, 8] > 0].min()), int(st[:, 9].max())
    out = {"gram": (g0, g1)}
    last_start = {"gram": int(st[:, 8].max())}
    for name in ("holdout", "loss"):
        b = (ctypes.c_ulonglong * 4)()
        getattr(lib, "omldm_tl_" + name)(b)
        out[name] = (int(b[0]), int(b[1]))
        last_start[name] = int(b[2])
    t0 = min(v[0] for v in out.values())
    for k, (a, b) in sorted(out.items(), key=lambda kv: kv[1][0]):
        print(f"  timeline {k:8s} start {(a - t0) / 1e3:6.1f} us  end {(b - t0) / 1e3:6.1f} us  ({(b - a) / 1e3:.1f} us); last block start {(last_start[k] - t0) / 1e3:6.1f} us")


dev = torch.device("cuda", 0)
B, d = 32768, 1000
pool = [R.generate_batch(R.KIND_REGRESSION, d, B, seed=5, batch_index=i, wire="f16s").to(dev) for i in range(3)]
for overlap, piped in (("true", "false"), ("true", "true"), ("false", "false")):
    jp = JobParameters(parallelism=1, jobName="t")
    jp.extra["pipelinedHoldout"] = piped
    eng = Engine(jp, 0, 1, dev, impl="fused")
    eng.handle_request({"id": 1, "request": "Create", "learner": {"name": "ORR", "hyperParameters": {"lambda": 1.0, "solveEvery": 1000000, "overlapLoss": overlap}},
                        "preProcessors": [], "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
    for i in range(30):
        eng.train_batch(pool[i % 3], d)
    eng.flush()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(40):
        eng.train_batch(pool[i % 3], d)
    eng.flush()
    e1.record()
    torch.cuda.synchronize()
    eng.flush()
    torch.cuda.synchronize()
    print(f"overlapLoss={overlap} pipelinedHoldout={piped}: {1000 * e0.elapsed_time(e1) / 40:.1f} us per step")
    for line in gram_phase_report():
        print(line)
    timeline()
    eng.close()
