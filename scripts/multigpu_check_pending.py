"""First multi-GPU run of the features written after round 1's GPU budget was spent (ROADMAP.md §1).  Run under
torchrun with >= 2 GPUs; prints 'PENDING_OK' on rank 0 when every check passed.

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/multigpu_check_pending.py

Checks: the scatter-fused Synchronous protocol (push kernel and GEMM-epilogue variants) against the NCCL baseline
engine, the FGM drain flush with unequal step counts per rank, and the timing of the three Synchronous variants."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from omldm_b200.engine import Engine
from omldm_b200.parallel.comm import PipelineComm
from omldm_b200.parallel.fused import make_driver
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ok = True


def check(name, cond):
    global ok
    t = torch.tensor([1 if cond else 0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(("PASS " if t.item() else "FAIL ") + name, flush=True)
    ok = ok and bool(t.item())


def run_orr(impl, tc_extra, d=992, B=2048, steps=12, timed=False):
    eng = Engine(JobParameters(parallelism=world, testSetSize=16), rank, world, dev, impl=impl)
    eng.handle_request({"id": 1, "request": "Create", "learner": {"name": "ORR", "hyperParameters": {"solveEvery": 4}},
                        "trainingConfiguration": dict({"protocol": "Synchronous", "nFeatures": d}, **tc_extra)})
    batches = [R.generate_batch(R.KIND_REGRESSION, d, B, seed=3, batch_index=world * i + rank).to(dev) for i in range(4)]
    for i in range(steps):
        eng.train_batch(batches[i % 4].clone(), d)
    ms = None
    if timed:
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        pool = [b.clone() for b in batches] * 8
        e0.record()
        for b in pool:
            eng.train_batch(b, d)
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / len(pool)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    out = eng.handle_request({"id": 1, "request": "Query", "requestId": 1})
    p = eng.pipelines[1]
    G = torch.from_numpy(p.learner.to_oracle().G).float().to(dev).reshape(-1)
    res = (json.loads(out[-1]) if out else None, G, type(p.driver).__name__, eng.comm_status(), ms)
    eng.close(); dist.barrier()
    return res


base = run_orr("nccl", {})
for label, tc in (("scatter (push kernel)", {"syncFusion": "scatter"}), ("gemm (TMA reduce into the owners)", {"syncFusion": "gemm"})):
    r = run_orr("fused", tc)
    if rank == 0:
        print("driver:", r[2], flush=True)
    check(f"ORR {label}: global Gram matrix == NCCL baseline",
          torch.allclose(r[1], base[1], rtol=2e-2, atol=2e-2 * float(base[1].abs().max())))
    check(f"ORR {label}: comm status 0", r[3] == 0)
    if rank == 0:
        check_local = r[0]["dataFitted"] == base[0]["dataFitted"]
    else:
        check_local = True
    check(f"ORR {label}: dataFitted", check_local)

# ---- FGM drain flush with unequal numbers of steps per rank
comm = PipelineComm(256, dev, rank, world)
e = torch.zeros(comm.n, device=dev)
drv = make_driver("FGM", comm, e, "mean", {"threshold": 1.0}, "fused")
for s in range(100 + 37 * rank):
    comm.buf.add_(0.001 * (rank + 1))
    drv.step()
drv.flush()
torch.cuda.synchronize(); dist.barrier()
rounds = torch.tensor([float(drv.device_rounds())], device=dev)
lo, hi = rounds.clone(), rounds.clone()
dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
check("fgm drain: every rank executed the same rounds, none left open, status 0",
      float(lo) == float(hi) >= 1 and comm.status() == 0 and drv.flush_timeouts == 0)
dist.barrier(); comm.close()

# ---- SingleLearner forwarding over peer memory vs the NCCL gather
res = {}
for transport in ("peer", "nccl"):
    jp = JobParameters(parallelism=world, testSetSize=16)
    jp.extra["forwardTransport"] = transport
    eng = Engine(jp, rank, world, dev, impl="fused")
    eng.handle_request({"id": 1, "request": "Create",
                        "learner": {"name": "HT", "hyperParameters": {"nClasses": 3, "gracePeriod": 100, "tieThreshold": 0.2}},
                        "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": 6}})
    for i in range(10):
        eng.train_batch(R.generate_batch(R.KIND_CLUSTERS, 6, 256, seed=5, batch_index=world * i + rank, n_aux=3).to(dev), 6)
    out = eng.handle_request({"id": 1, "request": "Query", "requestId": 2})
    res[transport] = (json.loads(out[-1]) if out else None, eng.comm_status())
    eng.close(); dist.barrier()
if rank == 0:
    same = res["peer"][0]["dataFitted"] == res["nccl"][0]["dataFitted"] == world * (10 * 256 - 16) \
        and abs(res["peer"][0]["score"] - res["nccl"][0]["score"]) < 0.05
else:
    same = True
check("single learner: peer-memory forwarding == NCCL gather (records fitted, score)", same)
check("single learner: comm status 0", res["peer"][1] == 0)

# ---- step time of the three Synchronous variants (ms per step, max over ranks)
for label, tc in (("two-barrier all-reduce", {}), ("scatter", {"syncFusion": "scatter"}), ("gemm", {"syncFusion": "gemm"})):
    r = run_orr("fused", tc, B=16384, steps=16, timed=True)
    if rank == 0:
        print(f"TIMING ORR d=992 B=16384 {label}: {1000 * r[4]:.1f} us/step", flush=True)

if rank == 0 and ok:
    print("PENDING_OK", flush=True)
torch.cuda.synchronize()
dist.barrier()
sys.stdout.flush()
os._exit(0)
