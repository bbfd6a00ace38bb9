"""Run under torchrun with >= 2 GPUs: fused protocol kernels over real NVLink peer memory vs
torch.distributed (NCCL) results.  Prints 'MULTIGPU_OK' on rank 0 when every check passed."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from omldm_b200.engine import Engine
from omldm_b200.parallel.comm import PipelineComm
from omldm_b200.parallel.fused import make_driver, query_reduce, device_barrier
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

for transport in ("ipc", "symm"):
    os.environ["OMLDM_USE_SYMM_MEM"] = "1" if transport == "symm" else "0"
    for n in (1000, 1 << 20):
        # ---- Synchronous, mean and sum-delta
        for mode_name in ("mean", "sum"):
            comm = PipelineComm(n, dev, rank, world)
            e = torch.zeros(comm.n, device=dev)
            torch.manual_seed(100 + rank)
            x = torch.randn(comm.n, device=dev)
            comm.buf.copy_(x)
            g0 = torch.randn(comm.n, device=dev, generator=torch.Generator(device=dev).manual_seed(5))
            if mode_name == "sum":
                e.copy_(g0)
            drv = make_driver("Synchronous", comm, e, mode_name, {}, "fused")
            ref = x.clone(); dist.all_reduce(ref)
            for _ in range(2):
                drv.step()
                if mode_name == "sum":
                    break
            torch.cuda.synchronize()
            if mode_name == "mean":
                check(f"{transport} sync mean n={n} mc={comm.multicast}", torch.allclose(comm.buf, ref / world, rtol=1e-5, atol=1e-5)
                      and torch.allclose(e, ref / world, rtol=1e-5, atol=1e-5))
            else:
                check(f"{transport} sync sum n={n}", torch.allclose(e, g0 + ref, rtol=1e-5, atol=1e-4) and float(comm.buf.abs().max()) == 0)
            check(f"{transport} status n={n}", comm.status() == 0)
            dist.barrier(); comm.close()
# ---- Asynchronous / SSP / EASGD / GM / FGM on IPC
os.environ["OMLDM_USE_SYMM_MEM"] = "0"
n = 4096
comm = PipelineComm(n, dev, rank, world)
e = torch.zeros(comm.n, device=dev)
drv = make_driver("Asynchronous", comm, e, "mean", {}, "fused")
upd = torch.full((comm.n,), float(rank + 1), device=dev)
comm.buf.add_(upd)
dist.barrier()
drv.step(); torch.cuda.synchronize(); dist.barrier()
lo, hi = comm.shard_bounds(rank)
expect = sum(range(1, world + 1)) / world
check("async hub shard holds the mean of all pushes", torch.allclose(comm.hub[lo:hi], torch.full((hi - lo,), expect, device=dev)))
check("async worker installed a hub value", torch.equal(comm.buf, e))
dist.barrier(); comm.close()

comm = PipelineComm(n, dev, rank, world)
e = torch.zeros(comm.n, device=dev)
drv = make_driver("SSP", comm, e, "mean", {"staleness": 1}, "fused")
for _ in range(5):
    comm.buf.add_(0.1)
    drv.step()
torch.cuda.synchronize(); dist.barrier()
check("ssp: no timeout, clocks advanced", comm.status() == 0)
dist.barrier(); comm.close()

comm = PipelineComm(n, dev, rank, world)
e = torch.zeros(comm.n, device=dev)
drv = make_driver("GM", comm, e, "mean", {"threshold": 10.0}, "fused")
comm.buf.fill_(0.001 * (rank + 1)); before = comm.buf.clone()
drv.step(); torch.cuda.synchronize()
check("gm: inside the ball nothing moves", torch.equal(comm.buf, before))
if rank == world - 1:
    comm.buf.fill_(1.0)
dist.barrier()
ref = comm.buf.clone(); dist.all_reduce(ref)
drv.step(); torch.cuda.synchronize()
check("gm: violation triggers the averaging round", torch.allclose(comm.buf, ref / world, rtol=1e-5))
dist.barrier(); comm.close()

comm = PipelineComm(256, dev, rank, world)
e = torch.zeros(comm.n, device=dev)
drv = make_driver("FGM", comm, e, "mean", {"threshold": 1.0}, "fused")
for s in range(150):
    comm.buf.add_(0.001 * (rank + 1))
    drv.step()
drv.flush()
torch.cuda.synchronize(); dist.barrier()
check("fgm: rounds happened and replicas agree", drv.device_rounds() >= 2 and comm.status() == 0)
dist.barrier(); comm.close()

comm = PipelineComm(16, dev, rank, world)
vals = torch.tensor([rank + 1.0, 2.0, 3.0, 4.0], device=dev)
out = query_reduce(comm, vals)
device_barrier(comm); torch.cuda.synchronize()
check("query reduce", torch.allclose(out, torch.tensor([sum(range(1, world + 1)), 2.0 * world, 3.0 * world, 4.0 * world], device=dev)))
dist.barrier(); comm.close()

# ---- engine end to end: fused == nccl baseline for Synchronous ORR (every synchronisation variant and both row
# formats) and PA
def run_engine(impl, name, hyper, kind, d, tc_extra=None, wire="f32", B=512, steps=10, extra=None, n_aux=0):
    jp = JobParameters(parallelism=world, testSetSize=16)
    jp.extra.update(extra or {})
    eng = Engine(jp, rank, world, dev, impl=impl)
    eng.handle_request({"id": 1, "request": "Create", "learner": {"name": name, "hyperParameters": hyper},
                        "trainingConfiguration": dict({"protocol": "Synchronous", "nFeatures": d}, **(tc_extra or {}))})
    for i in range(steps):
        eng.train_batch(R.generate_batch(kind, d, B, seed=3, batch_index=world * i + rank, wire=wire, n_aux=n_aux).to(dev), d)
    out = eng.handle_request({"id": 1, "request": "Query", "requestId": 1})
    p = eng.pipelines[1]
    if p.learner.merge_mode == "sum" and hasattr(p.learner, "to_oracle") and name == "ORR":
        flat = torch.from_numpy(p.learner.to_oracle().G).float().to(dev).reshape(-1)
    elif p.learner.merge_mode == "sum":
        flat = p.driver.e.clone()
    else:
        flat = p.comm.buf.clone()
    res = (json.loads(out[-1]) if out else None, flat, type(p.driver).__name__, eng.comm_status())
    eng.close(); dist.barrier()
    return res


cases = [("ORR fp32 rows, sharded hub (push kernel)", "ORR", {"solveEvery": 1}, R.KIND_REGRESSION, 100, {}, "f32"),
         ("ORR fp16 rows, sharded hub (push inside the Gram kernel)", "ORR", {"solveEvery": 2}, R.KIND_REGRESSION, 100, {}, "f16s"),
         ("ORR fp16 rows d=1000, sharded hub", "ORR", {"solveEvery": 4}, R.KIND_REGRESSION, 1000, {}, "f16s"),
         ("ORR fp32 rows, two-barrier all-reduce", "ORR", {"solveEvery": 1}, R.KIND_REGRESSION, 100, {"syncFusion": "none"}, "f32"),
         ("ORR fp16 rows, two-barrier all-reduce", "ORR", {"solveEvery": 1}, R.KIND_REGRESSION, 100, {"syncFusion": "none"}, "f16s"),
         ("ORR fp32 rows, scatter push", "ORR", {"solveEvery": 4}, R.KIND_REGRESSION, 992, {"syncFusion": "scatter"}, "f32"),
         ("ORR fp32 rows, scatter in the tf32 GEMM epilogue", "ORR", {"solveEvery": 4}, R.KIND_REGRESSION, 992, {"syncFusion": "gemm"}, "f32"),
         ("PA", "PA", {"C": 0.1}, R.KIND_BINARY, 28, {}, "f32")]
base_cache = {}
for label, name, hyper, kind, d, tc_extra, wire in cases:
    key = (name, d, wire, json.dumps(hyper))
    if key not in base_cache:
        base_cache[key] = run_engine("nccl", name, hyper, kind, d, {}, wire)
    base = base_cache[key]
    r = run_engine("fused", name, hyper, kind, d, tc_extra, wire)
    if rank == 0:
        print("drivers:", r[2], base[2], flush=True)
    check(f"engine {label}: fused replicas == nccl replicas",
          torch.allclose(r[1], base[1], rtol=2e-2, atol=2e-2 * float(base[1].abs().max())))
    check(f"engine {label}: comm status 0", r[3] == 0)
    if rank == 0:
        check_local = r[0]["dataFitted"] == base[0]["dataFitted"] == world * (10 * 512 - 16)
        if name == "ORR":
            check_local = check_local and abs(r[0]["score"] - base[0]["score"]) <= 0.05 * max(abs(base[0]["score"]), 1e-3)
    else:
        check_local = True
    check(f"engine {label}: dataFitted summed over workers (and the score for ORR)", check_local)

# ---- FGM rounds interleaved with queries through the engine (round counter and query counter are independent):
# few queries / many rounds, and many queries / few rounds
for label, q_every, thr in (("rounds > queries", 25, 0.05), ("queries > rounds", 1, 50.0)):
    eng = Engine(JobParameters(parallelism=world, testSetSize=16), rank, world, dev, impl="fused")
    eng.handle_request({"id": 1, "request": "Create", "learner": {"name": "PA", "hyperParameters": {"C": 0.5}},
                        "trainingConfiguration": {"protocol": "FGM", "nFeatures": 28, "threshold": thr}})
    fitted_ok, n_q = True, 0
    for i in range(60):
        eng.train_batch(R.generate_batch(R.KIND_BINARY, 28, 256, seed=11, batch_index=world * i + rank).to(dev), 28)
        if (i + 1) % q_every == 0:
            out = eng.handle_request({"id": 1, "request": "Query", "requestId": 100 + i})
            n_q += 1
            if rank == 0:
                q = json.loads(out[-1])
                seen = (i + 1) * 256
                fitted_ok = fitted_ok and q["dataFitted"] == world * (seen - 16)
    eng.flush()
    rounds = eng.pipelines[1].driver.device_rounds()
    st = eng.comm_status()
    if rank == 0:
        print(f"fgm + query ({label}): {rounds} rounds, {n_q} queries", flush=True)
    check(f"fgm + query ({label}): every merged dataFitted exact, comm status 0", fitted_ok and st == 0)
    eng.close(); dist.barrier()

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
lo_r, hi_r = rounds.clone(), rounds.clone()
dist.all_reduce(lo_r, op=dist.ReduceOp.MIN); dist.all_reduce(hi_r, op=dist.ReduceOp.MAX)
check("fgm drain: every rank executed the same rounds, none left open, status 0",
      float(lo_r) == float(hi_r) >= 1 and comm.status() == 0 and drv.flush_timeouts == 0)
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

if rank == 0 and ok:
    print("MULTIGPU_OK", flush=True)
torch.cuda.synchronize()
dist.barrier()
sys.stdout.flush()
os._exit(0)   # skip interpreter teardown of peer-mapped allocations (has been seen to stall)
