"""Job entry point: `python -m omldm_b200.cli --parallelism 8 ...` (under torchrun for > 1 GPU).

Mirrors `omldm.Job.main` (Job.scala:110-168): `--key value` flags with the reference's names and
defaults (utils/DefaultJobParameters.scala), requests / training / forecasting streams in, responses /
predictions / performance out.  Kafka topics become files or synthetic generators:

  --requests FILE            newline-delimited JSON Requests (the `requests` topic)
  --trainingData SPEC        FILE of JSON DataInstances, or  synthetic:<kind>:<d>[:<n_aux>]
                             (kind = binary | regression | clusters | multiclass)
  --forecastingData FILE     JSON DataInstances with operation "forecasting"
  --batch N --batches M      mini-batch size per worker and number of mini-batches
  --queryEvery K             issue the Query requests of the request file every K batches
  --outDir DIR               responses / predictions / performance are written there (one file per topic name)
  --engine gpu|host          gpu = one worker + hub shard per GPU (default); host = fp64 message-driven oracle
  --impl fused|nccl          fused protocol kernels (default) or the torch.distributed baseline
  --checkpointing true --checkInterval MS --stateBackend DIR [--resume true]
  --injectStall rank:ms      fault injection: stall one rank's host loop before a step
"""
from __future__ import annotations

import json
import os
import sys
import time
from typing import List

from .utils.config import JobParameters

KINDS = {"binary": 0, "regression": 1, "clusters": 2, "multiclass": 3}


def _read_lines(path: str) -> List[str]:
    if not path or not os.path.exists(path):
        return []
    with open(path) as f:
        return [l for l in (x.strip() for x in f) if l]


def _write(out_dir: str, topic: str, lines: List[str]) -> None:
    if not lines:
        return
    os.makedirs(out_dir, exist_ok=True)
    with open(os.path.join(out_dir, topic + ".jsonl"), "a") as f:
        for l in lines:
            f.write(l + "\n")


def run_host(params: JobParameters) -> int:
    from .runtime.job import HostJob
    job = HostJob(params)
    ex = params.extra
    for r in _read_lines(ex.get("requests", "")):
        if json.loads(r).get("request") != "Query":
            job.submit_request(r)
    queries = [r for r in _read_lines(ex.get("requests", "")) if json.loads(r).get("request") == "Query"]
    for rec in _read_lines(ex.get("trainingData", "")):
        job.submit_data(rec)
    for rec in _read_lines(ex.get("forecastingData", "")):
        job.submit_data(rec)
    for q in queries:
        job.submit_request(q)
    js = job.run()
    out = ex.get("outDir", "out")
    _write(out, params.topics["responses"], job.responses)
    _write(out, params.topics["predictions"], job.predictions)
    _write(out, params.topics["performance"], job.performance)
    print(job.performance[-1] if job.performance else "{}")
    return 0


def run_gpu(params: JobParameters) -> int:
    import torch
    import torch.distributed as dist
    from .engine import Engine
    from .ops import functional as F
    from .stream import ring as R
    from .utils.checkpoint import PeriodicCheckpointer, load_checkpoint
    from .utils.numa import bind_to_gpu_numa

    ex = params.extra
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(local)
        bind_to_gpu_numa(local)
    dev = torch.device("cuda", local) if use_cuda else torch.device("cpu")
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl" if use_cuda else "gloo")
    params.parallelism = world
    impl = ex.get("impl", "fused")
    if impl == "nccl":
        F.set_impl("torch")
    eng = Engine(params, rank, world, dev, impl=impl)
    reqs = [json.loads(r) for r in _read_lines(ex.get("requests", ""))]
    queries = [r for r in reqs if r.get("request") == "Query"]
    for r in reqs:
        if r.get("request") != "Query":
            eng.handle_request(r)
    B = int(ex.get("batch", 4096))
    n_batches = int(ex.get("batches", 100))
    query_every = int(ex.get("queryEvery", 0))
    spec = ex.get("trainingData", "synthetic:binary:28")
    stall_rank, stall_ms = -1, 0
    if params.injectStall:
        a, b = params.injectStall.split(":")
        stall_rank, stall_ms = int(a), int(b)
    ckpt = None
    if params.checkpointing:
        ckpt = PeriodicCheckpointer(eng, params.stateBackend, params.checkInterval)
        if str(ex.get("resume", "false")).lower() in ("1", "true"):
            step0 = load_checkpoint(eng, params.stateBackend)
            if rank == 0:
                print(f"resumed from step {step0}", file=sys.stderr)
    responses: List[str] = []
    if spec.startswith("synthetic:"):
        parts = spec.split(":")
        kind, d = KINDS[parts[1]], int(parts[2])
        n_aux = int(parts[3]) if len(parts) > 3 else 0
        topic = R.Topic(d, B, n_slots=4)
        topic.start_synthetic(kind, seed=1000 + rank, n_aux=n_aux, max_batches=n_batches, n_threads=2)
        step = 0
        while True:
            got = topic.get(30000)
            if got is None:
                break
            slot, Zh, nrec = got
            if rank == stall_rank and step == n_batches // 2:
                time.sleep(stall_ms / 1000.0)
            if use_cuda:
                eng.train_batch_host(Zh, d)
                if eng.last_copy_event is not None:
                    eng.last_copy_event.synchronize()
            else:
                eng.train_batch(Zh[:nrec].clone(), d)
            topic.release(slot)
            step += 1
            if query_every and step % query_every == 0:
                for q in queries:
                    responses += eng.handle_request(dict(q, requestId=int(q.get("requestId", 0)) + step))
            if ckpt:
                ckpt.maybe_save(step)
        topic.stop()
    else:
        # JSON records from a file: every rank takes its share (record i goes to worker i % world)
        lines = [l for l in _read_lines(spec) if l.startswith("{")][rank::world]
        text = ("\n".join(lines) + "\n").encode()
        d = len(json.loads(lines[0])["numericalFeatures"]) + len(json.loads(lines[0]).get("discreteFeatures") or [])
        off = 0
        while off < len(text):
            rows, nf, used = R.parse_json_records(text, d, B, offset=off)
            off += max(used, 1)
            if rows.shape[0]:
                eng.train_batch(rows[rows[:, d] == 1.0].to(dev), d)
    fc = [l for l in _read_lines(ex.get("forecastingData", "")) if l.startswith("{")][rank::world]
    preds: List[str] = []
    if fc and eng.d is not None:
        from .api.pojos import DataInstance
        rows, nf, _ = R.parse_json_records(("\n".join(fc) + "\n").encode(), eng.d, len(fc))
        rows[:, eng.d] = 1.0
        preds = eng.forecast_batch(rows.to(dev), [DataInstance.from_json(l) for l in fc][: rows.shape[0]])
    for q in queries:
        responses += eng.handle_request(q)
    js = eng.terminate()
    out = ex.get("outDir", "out")
    if rank == 0:
        _write(out, params.topics["responses"], responses)
        _write(out, params.topics["performance"], eng.performance)
        print(eng.performance[-1] if eng.performance else "{}")
    _write(out, params.topics["predictions"] + (f".rank{rank}" if world > 1 else ""), preds)
    status = eng.comm_status()
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    return 0 if status == 0 else 3


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if "--help" in argv or "-h" in argv:
        print(__doc__)
        return 0
    params = JobParameters.from_args(argv)
    if params.engine == "host":
        return run_host(params)
    return run_gpu(params)


if __name__ == "__main__":
    sys.exit(main())
