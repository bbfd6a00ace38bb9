"""Profiling target: a few eager ORR steps (graphs off) + an emulated 8-rank fused sync, so that
ncu sees every kernel of the flagship step as an ordinary launch."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from omldm_b200.engine import Engine
from omldm_b200.ops import _lib
from omldm_b200.parallel.comm import PipelineComm
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters

dev = torch.device("cuda", 0)
d, B = 1000, 16384
p = JobParameters(parallelism=1); p.extra["graphs"] = "false"
eng = Engine(p, 0, 1, dev)
eng.handle_request({"id": 1, "request": "Create", "learner": {"name": "ORR", "hyperParameters": {"solveEvery": 1000}},
                    "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
pool = [R.generate_batch(R.KIND_REGRESSION, d, B, seed=1, batch_index=i).to(dev) for i in range(4)]
for i in range(8):
    eng.train_batch(pool[i % 4], d)
torch.cuda.synchronize()
# fused synchronous step on 8 emulated ranks, 4 MB model (the cfg-2 sync shape)
k, n = 8, 1024 * 1024
comm = PipelineComm(n, dev, 0, 1, None, emulate=k)
e = torch.zeros(k, comm.n, device=dev)
for r in range(k):
    comm.bufs[r].normal_()
for _ in range(3):
    _lib.check(_lib.kernels().omldm_sync_allreduce(comm.struct_ptr(), _lib.ptr(e), ctypes.c_size_t(comm.n),
               ctypes.c_float(1.0), ctypes.c_int(1), ctypes.c_uint32(0), ctypes.c_int(0), _lib.stream_ptr()), "sync")
torch.cuda.synchronize()
print("done", eng.pipelines[1].learner.counters())
