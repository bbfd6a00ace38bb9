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
                             (a request with an extra key "atBatch": k is applied at the boundary before step k: Create /
                             Update / Query / Delete in mid-stream; forecasts are served in slices between batches)
  --wire f32|f16|f16s        row format of the synthetic topic (f16s: fp16 staged rows, trained on in place by ORR)
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
    from .api.pojos import DataInstance
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
    B = int(ex.get("batch", 4096))
    n_batches = int(ex.get("batches", 100))
    query_every = int(ex.get("queryEvery", 0))
    spec = ex.get("trainingData", "synthetic:binary:28")
    wire = ex.get("wire", "f32")
    stall_rank, stall_ms = -1, 0
    if params.injectStall:
        a, b = params.injectStall.split(":")
        stall_rank, stall_ms = int(a), int(b)

    # ---- the `requests` topic.  A request may carry "atBatch": k - it is then applied at the boundary before step k
    # (requests in mid-stream: Create / Update / Delete while data flows, PipelineMap.scala:31-57); without it, Create /
    # Update / Delete come before the first record and Query at the end (and every --queryEvery steps).
    reqs = [json.loads(r) for r in _read_lines(ex.get("requests", ""))]
    timed = sorted([r for r in reqs if "atBatch" in r], key=lambda r: int(r["atBatch"]))
    queries = [r for r in reqs if r.get("request") == "Query" and "atBatch" not in r]
    start_reqs = [r for r in reqs if r.get("request") != "Query" and "atBatch" not in r]
    responses: List[str] = []

    step0 = 0
    ckpt = None
    resumed = False
    if params.checkpointing:
        ckpt = PeriodicCheckpointer(eng, params.stateBackend, params.checkInterval)
        if str(ex.get("resume", "false")).lower() in ("1", "true"):
            got = load_checkpoint(eng, params.stateBackend)
            if got is not None:
                step0, resumed = int(got), True       # the stream position: nothing before it is trained on again
                if rank == 0:
                    print(f"resumed from step {step0}", file=sys.stderr)
    if not resumed:
        for r in start_reqs:
            eng.handle_request(r)
    else:
        timed = [r for r in timed if int(r["atBatch"]) >= step0]

    # ---- the `forecastingData` topic: this worker's partition, served in slices between training batches
    fc = [l for l in _read_lines(ex.get("forecastingData", "")) if l.startswith("{")][rank::world]
    preds: List[str] = []
    fc_pos = 0

    def serve_forecasts(upto: int) -> None:
        nonlocal fc_pos
        if fc_pos >= upto or eng.d is None or not eng.pipelines:
            return
        chunk = fc[fc_pos:upto]
        rows, _, _ = R.parse_json_records(("\n".join(chunk) + "\n").encode(), eng.d, len(chunk))
        if rows.shape[0]:
            rows[:, eng.d] = 1.0
            preds.extend(eng.forecast_batch(rows.to(dev), [DataInstance.from_json(l) for l in chunk][: rows.shape[0]]))
        fc_pos = upto

    def boundary(step: int, total: int) -> None:
        """Step boundary: requests due now, forecasts that arrived meanwhile, periodic queries, checkpoints."""
        nonlocal responses
        while timed and int(timed[0]["atBatch"]) <= step:
            r = dict(timed.pop(0))
            r.pop("atBatch")
            out = eng.handle_request(r)
            responses += out
        if fc and total > 0:
            serve_forecasts(min(len(fc), (step * len(fc)) // total))
        if query_every and step and step % query_every == 0:
            for q in queries:
                responses += eng.handle_request(dict(q, requestId=int(q.get("requestId", 0)) + step))
        if ckpt and step > step0:
            ckpt.maybe_save(step)

    if spec.startswith("synthetic:"):
        parts = spec.split(":")
        kind, d = KINDS[parts[1]], int(parts[2])
        n_aux = int(parts[3]) if len(parts) > 3 else 0
        topic = R.Topic(d, B, n_slots=4, wire=wire)
        topic.start_synthetic(kind, seed=1000 + rank, n_aux=n_aux, max_batches=n_batches, n_threads=2, first_batch=step0)
        step = step0
        pending = []
        while step < n_batches:
            boundary(step, n_batches)
            got = topic.get(30000)
            if got is None:
                break
            slot, Zh, nrec = got
            if rank == stall_rank and step == n_batches // 2:
                time.sleep(stall_ms / 1000.0)
            if use_cuda:
                eng.train_batch_host(Zh, d, wire=wire)
                pending.append((slot, eng.last_copy_event))
                while len(pending) > 1:              # a slot goes back once its H2D copy has completed
                    s0, ev = pending.pop(0)
                    if ev is not None:
                        ev.synchronize()
                    topic.release(s0)
            else:
                eng.train_batch(Zh[:nrec].clone(), d)
                topic.release(slot)
            step += 1
        for s0, ev in pending:
            if ev is not None:
                ev.synchronize()
            topic.release(s0)
        topic.stop()
        boundary(step, n_batches)
    else:
        # JSON records from a file: record i goes to worker i % world.  The step schedule is common to all ranks -
        # the same number of steps, the same batch shape, short or missing batches padded with empty rows - because
        # the barrier protocols need every rank in every step.
        lines = [l for l in _read_lines(spec) if l.startswith("{")]
        first = json.loads(lines[0])
        d = len(first["numericalFeatures"]) + len(first.get("discreteFeatures") or [])
        mine = lines[rank::world]
        per_rank_max = (len(lines) + world - 1) // world
        n_steps = (per_rank_max + B - 1) // B
        ld = F.pad32(d + 2)
        for step in range(step0, n_steps):
            boundary(step, n_steps)
            chunk = mine[step * B:(step + 1) * B]
            Z = torch.zeros(B, ld)
            if chunk:
                rows, _, _ = R.parse_json_records(("\n".join(chunk) + "\n").encode(), d, len(chunk))
                rows = rows[rows[:, d] == 1.0]
                Z[: rows.shape[0]] = rows
            eng.train_batch(Z.to(dev), d)
        boundary(n_steps, n_steps)
    serve_forecasts(len(fc))
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
