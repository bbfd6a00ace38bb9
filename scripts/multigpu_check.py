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

# ---- engine end to end: fused == nccl baseline for Synchronous ORR and PA
for name, hyper, kind, d in (("ORR", {"solveEvery": 1}, R.KIND_REGRESSION, 100), ("PA", {"C": 0.1}, R.KIND_BINARY, 28)):
    res = {}
    for impl in ("fused", "nccl"):
        eng = Engine(JobParameters(parallelism=world, testSetSize=16), rank, world, dev, impl=impl)
        eng.handle_request({"id": 1, "request": "Create", "learner": {"name": name, "hyperParameters": hyper},
                            "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
        for i in range(10):
            eng.train_batch(R.generate_batch(kind, d, 512, seed=3, batch_index=world * i + rank).to(dev), d)
        out = eng.handle_request({"id": 1, "request": "Query", "requestId": 1})
        res[impl] = json.loads(out[-1]) if out else None
        p = eng.pipelines[1]
        if p.learner.merge_mode == "sum":
            res[impl + "_flat"] = torch.from_numpy(p.learner.to_oracle().G).float().to(dev).reshape(-1)
        else:
            res[impl + "_flat"] = p.comm.buf.clone()
        res[impl + "_driver"] = type(p.driver).__name__
        eng.close(); dist.barrier()
    if rank == 0:
        print("drivers:", res["fused_driver"], res["nccl_driver"], flush=True)
    check(f"engine {name}: fused replicas == nccl replicas",
          torch.allclose(res["fused_flat"], res["nccl_flat"], rtol=2e-2, atol=2e-2 * float(res["nccl_flat"].abs().max())))
    if rank == 0:
        check_local = res["fused"]["dataFitted"] == res["nccl"]["dataFitted"] == world * (10 * 512 - 16)
    else:
        check_local = True
    check(f"engine {name}: dataFitted summed over workers", check_local)
if rank == 0 and ok:
    print("MULTIGPU_OK", flush=True)
torch.cuda.synchronize()
dist.barrier()
sys.stdout.flush()
os._exit(0)   # skip interpreter teardown of peer-mapped allocations (has been seen to stall)
