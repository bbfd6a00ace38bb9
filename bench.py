#!/usr/bin/env python
"""Headline benchmark: records/sec of distributed online learning on N B200s of one node.

Default config = BASELINE.json config 2 (online ridge regression, 1000-dim synthetic stream,
Synchronous protocol); `--config` selects the other BASELINE configs:
  orr     ORR 1k-d, Synchronous                          (cfg 2, default)
  pa      PA binary classifier 28-d, Asynchronous         (cfg 1)
  fgm     RegressorPA + PolynomialFeatures + StandardScaler, FGM (cfg 3)
  kmeans  K-means K=256 on 128-d, SSP                     (cfg 4)
  mcpa    MultiClassPA 100k-d sparse->dense, Synchronous  (cfg 5)

One step = every GPU ingests one mini-batch, updates every pipeline and runs the protocol's fused
synchronisation kernel (weak scaling: per-GPU batch is fixed).  Prints ONE JSON line on rank 0.

  value  : device-timed (CUDA events, max over ranks) aggregate records/s, inputs resident on the
           device and rotated over a set larger than L2.  The window of exactly --steps steps is repeated
           (`windows`) until at least ~0.4 s of device time has been measured; value / ms_per_step are the
           MEDIAN window, every window time is in `windows_ms` - a 4 ms window cannot decide the headline,
           and the nvidia-smi clock samples fall inside the timed windows
  e2e    : same metric through Engine.train_batch_host: pinned host ring slot -> async H2D -> train
           -> loss copied back to pinned host memory every step (median of 5 windows)
  vs_baseline : value / the same measurement of the in-repo PyTorch + NCCL (+cuBLAS) baseline, run in the
           same invocation (`baseline` block).  It is NOT the reference: the reference cannot run here
            and publishes no numbers
  sync_check : (more than one GPU) a short replay whose global statistics are compared across ranks and
           against an NCCL all-reduce of the per-rank contributions

`--impl nccl` runs the in-repo PyTorch + NCCL (+cuBLAS) baseline; `--impl reference` reports that
the reference (a Flink/Kafka/Scala job) cannot run here.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

CONFIGS = {
    "orr": dict(model="ORR", d=1000, batch=32768, protocol="Synchronous", kind=1, n_aux=0, wire="f16s",
                hyper={"lambda": 1.0, "solveEvery": 16}, tc={}, pre=[],
                name="ORR 1k-dim synthetic stream, synchronous protocol"),
    "pa": dict(model="PA", d=28, batch=65536, protocol="Asynchronous", kind=0, n_aux=0,
               hyper={"C": 0.01, "trainingMode": "minibatch"}, tc={}, pre=[],
               name="PA binary classifier, 28-dim dense synthetic stream, async protocol"),
    "fgm": dict(model="RegressorPA", d=28, batch=16384, protocol="FGM", kind=1, n_aux=0,
                hyper={"C": 0.01, "epsilon": 0.1, "trainingMode": "minibatch"}, tc={"threshold": 0.05},
                pre=[{"name": "PolynomialFeatures", "hyperParameters": {"degree": 2}}, {"name": "StandardScaler"}],
                name="PA regressor + polynomial features + standard scaler, FGM protocol"),
    "kmeans": dict(model="K-means", d=128, batch=16384, protocol="SSP", kind=2, n_aux=256,
                   hyper={"k": 256}, tc={"staleness": 3}, pre=[],
                   name="Online k-means k=256 on 128-dim synthetic stream, SSP protocol"),
    "mcpa": dict(model="MultiClassPA", d=100000, batch=16384, protocol="Synchronous", kind=4, n_aux=10, nnz=128,
                 hyper={"nClasses": 10, "C": 0.1}, tc={}, pre=[],
                 name="Multiclass PA 100k-dim sparse-to-dense stream, hub sharded over all GPUs"),
}


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.proc = None
        self.lines = []
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for nme, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "fused", "nccl", "reference"])
    ap.add_argument("--config", default="orr", choices=sorted(CONFIGS))
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-baseline", action="store_true",
                    help="skip the in-repo PyTorch + NCCL baseline arm (then vs_baseline is null)")
    ap.add_argument("--pipelined-holdout", action="store_true",
                    help="run the holdout exchange of the next batch on its own stream behind the current training "
                         "kernels (Engine option `pipelinedHoldout`) instead of inside the step's own graph")
    ap.add_argument("--min-timed-ms", type=float, default=400.0,
                    help="repeat the --steps window until this much device time has been measured")
    ap.add_argument("--wire", default="auto", choices=["auto", "f32", "f16", "f16s"],
                    help="row format of the records (pinned host ring and, for f16s, the device-resident batches): "
                         "f32 staged rows, f16 compact wire rows expanded on the GPU, f16s fp16 staged rows consumed "
                         "in place by the fp16 Gram kernel.  auto = f16s for ORR, f16 for the other dense configs")
    ap.add_argument("--sparse-direct", action="store_true", default=None,
                    help="sparse configs: train straight from the sparse rows (default)")
    ap.add_argument("--densify", dest="sparse_direct", action="store_false",
                    help="sparse configs: rebuild the dense batch every step and run the two GEMMs over it")
    ap.add_argument("--dry-run-cpu", action="store_true",
                    help="exercise this script on the CPU path (no GPU, wall-clock timing, tiny shapes); the line it "
                         "prints is marked dry_run and is not a measurement")
    ap.add_argument("--tc", action="append", default=[], help="extra trainingConfiguration key=value")
    ap.add_argument("--hyper", action="append", default=[], help="extra learner hyper-parameter key=value")
    args = ap.parse_args()

    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:        # under torchrun only rank 0 prints the line
            return
        print(json.dumps({"impl": "reference", "unavailable":
                          "reference is a Scala/Flink/Kafka Maven job with 3 out-of-tree SNAPSHOT dependencies; "
                          "no JVM/Maven/Kafka in this image and it is not pip-installable (no setup.py/pyproject)"}))
        return

    # stdout carries the ONE JSON line and nothing else: whatever a library prints to file descriptor 1 meanwhile (NCCL's
    # version banner does) goes to stderr; the line itself is written to the saved descriptor at the end
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    from omldm_b200.engine import Engine
    from omldm_b200.ops import _lib
    from omldm_b200.ops import functional as F
    from omldm_b200.stream import ring as R
    from omldm_b200.utils.config import JobParameters

    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # stdout carries the ONE JSON line only
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dry = args.dry_run_cpu
    if not dry and not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device", "metric": "records/sec"}), file=real_stdout, flush=True)
        return
    if dry:
        dev, numa = torch.device("cpu"), None
        if world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("gloo")
    else:
        torch.cuda.set_device(local)
        from omldm_b200.utils.numa import bind_to_gpu_numa
        numa = bind_to_gpu_numa(local)      # pinned ring + producer threads on the GPU's socket
        dev = torch.device("cuda", local)
        if world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=dev)

    class _WallEvent:                       # stand-in for torch.cuda.Event in the CPU dry run
        def record(self):
            self.t = time.perf_counter()

        def elapsed_time(self, other):
            return 1000.0 * (other.t - self.t)

    def dev_sync():
        if not dry:
            torch.cuda.synchronize()

    def sync_all() -> None:
        if world > 1:
            dist.barrier()
        dev_sync()

    cfg0 = CONFIGS[args.config]
    B = args.batch or cfg0["batch"]
    d = cfg0["d"]
    sparse = cfg0["kind"] == 4
    nnz = cfg0.get("nnz", 0)
    wire = args.wire
    if wire == "auto":
        wire = cfg0.get("wire", "f16") if not sparse else "f32"
    if sparse:
        wire = "f32"
    sparse_direct = bool(sparse and (args.sparse_direct is None or args.sparse_direct))
    dev_wire = "f16s" if wire == "f16s" else "f32"      # format of the device-resident batches of the timed loop
    tc = dict(cfg0["tc"], protocol=cfg0["protocol"], nFeatures=d)
    tc.update(kv.split("=", 1) for kv in args.tc)
    cfg = dict(cfg0, hyper=dict(cfg0["hyper"], **dict(kv.split("=", 1) for kv in args.hyper)))
    row_bytes = (F.f16s_pitch(d) * 2) if dev_wire == "f16s" else F.pad32(d + 2) * 4

    # ---- inputs: a set of distinct batches larger than L2 (126 MB), resident on the device
    n_rot = max(3, min(64, -(-160_000_000 // (B * row_bytes)) + 1))
    if sparse:      # device-resident sparse slots
        n_rot = max(4, min(2048, -(-160_000_000 // (B * (2 * nnz + 1) * 4)))) if sparse_direct else 4
        host_pool = [R.generate_sparse_batch(d, B, nnz, n_aux=cfg["n_aux"], seed=1234 + rank, batch_index=i, pin=not dry)
                     for i in range(n_rot)]
    else:
        host_pool = [R.generate_batch(cfg["kind"], d, B, seed=1234 + rank, batch_index=i, n_aux=cfg["n_aux"], pin=not dry,
                                      wire=dev_wire) for i in range(n_rot)]
    dev_pool = [h.to(dev) for h in host_pool]

    def make_engine(impl: str):
        if impl == "nccl":
            F.set_impl("torch")
            if not dry:
                torch.backends.cuda.matmul.allow_tf32 = True     # same math class as the tcgen05 tf32 path
        else:
            F.set_impl("auto")
        params = JobParameters(parallelism=world, jobName="bench_" + args.config)
        params.extra["sparseDirect"] = "true" if sparse_direct else "false"
        if impl != "nccl" and args.pipelined_holdout:
            # the holdout exchange of the next batch runs behind the current training kernels (Engine option; measured
            # on a B200: 41.2 against 43.0 us per ORR step with 3 rotating batches, 50.3 against 48.5 us with the 4 of
            # this script - no clear gain, so it stays off)
            params.extra["pipelinedHoldout"] = "true"
        if impl == "nccl":      # the in-repo baseline: torch.distributed collectives + stock PyTorch ops
            from omldm_b200 import baseline
            eng = baseline.make_engine(params, rank, world, dev)
        else:
            eng = Engine(params, rank, world, dev, impl=impl)
        eng.handle_request({"id": 1, "request": "Create",
                            "learner": {"name": cfg["model"], "hyperParameters": cfg["hyper"]},
                            "preProcessors": cfg["pre"], "trainingConfiguration": tc})
        return eng

    def measure(impl: str, min_ms: float, with_e2e: bool, sample_clocks: bool) -> dict:
        eng = make_engine(impl)

        def step_device(i: int) -> None:
            if sparse:
                eng.train_sparse(dev_pool[i % n_rot], B, nnz, d)
            else:
                eng.train_batch(dev_pool[i % n_rot], d)

        # untimed set-up: every batch buffer gets its CUDA graph captured
        n_keys = n_rot * (3 if str(tc.get("syncFusion", "")).lower() == "gemm" else 1)
        for i in range(3 * n_keys + 1):
            step_device(i)
        for i in range(max(args.warmup, 3)):
            step_device(i)
        eng.flush()
        sync_all()
        sampler = ClockSampler(local) if (sample_clocks and rank == 0 and not dry) else None
        mk = (lambda: _WallEvent()) if dry else (lambda: torch.cuda.Event(enable_timing=True))
        windows, launches, it = [], 0, args.warmup
        prof_window = (not dry) and impl != "nccl" and os.environ.get("OMLDM_PROFILE_WINDOW") == "1"
        total_ms, max_windows = 0.0, 4000
        while True:
            sync_all()
            _lib.reset_launch_count()
            if prof_window:      # `ncu --profile-from-start off`: capture exactly this window (never a bench value)
                torch.cuda.cudart().cudaProfilerStart()
            e0, e1 = mk(), mk()
            e0.record()
            for i in range(args.steps):
                step_device(it + i)
            eng.flush()          # inside the timed region: the protocol's end-of-stream settlement is part of the work
            e1.record()
            dev_sync()
            if prof_window:
                torch.cuda.cudart().cudaProfilerStop()
            it += args.steps
            launches = _lib.launch_count()
            ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)      # max over ranks: also makes every rank stop together
            windows.append(float(ms.item()))
            total_ms += windows[-1]
            if total_ms >= min_ms or len(windows) >= max_windows or dry:
                break
        clocks = sampler.stop() if sampler else None
        srt = sorted(windows)
        med = srt[len(srt) // 2]
        out = {"value": world * B * args.steps / (med / 1e3), "ms_per_step": med / args.steps, "windows": len(windows),
               "windows_ms": [round(w, 4) for w in windows[:64]], "window_ms_min": srt[0], "window_ms_max": srt[-1],
               "launches": launches, "clocks": clocks, "e2e": None}

        # ---- end to end: pinned ring slot -> H2D -> train -> loss D2H, every step
        if with_e2e:
            try:
                if sparse:
                    topic = R.SparseTopic(d, B, nnz, n_slots=8)
                    topic.start_synthetic(seed=99 + rank, n_aux=cfg["n_aux"], mode=R.MODE_REPLAY, n_threads=2)
                else:
                    topic = R.Topic(d, B, n_slots=max(4, min(n_rot, 8)), wire=wire)
                    topic.start_synthetic(cfg["kind"], seed=99 + rank, n_aux=cfg["n_aux"], mode=R.MODE_REPLAY,
                                          n_threads=2)
                pending = []   # (slot, event recorded after the slot's H2D copy)

                def step_e2e():
                    slot, Zh, _ = topic.get(10000)
                    loss = eng.train_sparse_host(Zh, B, nnz, d) if sparse else eng.train_batch_host(Zh, d, wire=wire)
                    pending.append((slot, eng.last_copy_event))
                    while len(pending) > 1:          # give a slot back once its H2D copy has completed
                        s0, ev = pending.pop(0)
                        if ev is not None:
                            ev.synchronize()
                        topic.release(s0)
                    return loss
                for _ in range(max(8, args.warmup)):   # includes graph capture for the two staging buffers
                    step_e2e()
                eng.flush()
                # the box's own ceiling for this path: every rank copies a pinned slot to its GPU at the same time,
                # nothing else running (one cudaMemcpyAsync per slot, like the engine)
                ceiling = None
                if not dry:
                    src, dst = topic.storage[0], torch.empty_like(topic.storage[0], device=dev)
                    dst.copy_(src, non_blocking=True)
                    sync_all()
                    n_cp = max(10, int(1.0e9 // topic.slot_bytes))       # ~1 GB per pass: long enough to measure bandwidth
                    ceiling = 0.0
                    for _ in range(3):                                   # best of three passes: it is meant to be a ceiling
                        sync_all()
                        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        c0.record()
                        for _ in range(n_cp):
                            dst.copy_(src, non_blocking=True)
                        c1.record()
                        dev_sync()
                        cms = torch.tensor([c0.elapsed_time(c1)], device=dev)
                        if world > 1:
                            dist.all_reduce(cms, op=dist.ReduceOp.MAX)
                        ceiling = max(ceiling, n_cp * topic.slot_bytes / (float(cms.item()) / 1e3) / 1e9)
                    del dst
                e_windows = []
                for _ in range(1 if dry else 5):
                    sync_all()
                    t0 = time.perf_counter()
                    for _ in range(args.steps):
                        loss_host = step_e2e()
                    eng.flush()
                    dev_sync()
                    dt = torch.tensor([time.perf_counter() - t0], device=dev)
                    if world > 1:
                        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
                    e_windows.append(float(dt.item()))
                e_med = sorted(e_windows)[len(e_windows) // 2]
                h2d = topic.slot_bytes
                out["e2e"] = {"value": world * B * args.steps / e_med, "unit": "records/s",
                              "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 32,
                              "wire": "sparse ELL (int32 index + fp32 value)" if sparse else wire,
                              "windows_s": [round(w, 5) for w in e_windows],
                              "h2d_gbs_per_gpu": h2d * args.steps / e_med / 1e9,
                              "h2d_ceiling_gbs_per_gpu": ceiling,
                              "fraction_of_h2d_ceiling": (h2d * args.steps / e_med / 1e9 / ceiling) if ceiling else None,
                              "h2d_ceiling_note": "pinned -> device cudaMemcpyAsync of one slot, all ranks copying at "
                                                  "once, nothing else running (max over ranks, best of 3 passes)",
                              "last_loss_sum": float(loss_host[0])}
                topic.stop()
            except Exception as ex:      # the device-timed line above stays valid; report why e2e is missing
                if world > 1:
                    raise
                out["e2e"] = {"error": f"{type(ex).__name__}: {ex}"}
        out["status"] = eng.comm_status()
        p = eng.pipelines[1]
        out["protocol"] = p.protocol
        out["sync_every"] = p.driver.sync_every
        out["model_bytes"] = getattr(p.driver, "n", p.comm.n) * 4
        out["driver"] = type(p.driver).__name__
        out["transport"] = "NVLS multimem" if getattr(p.comm, "multicast", False) else "P2P ld/st + TMA reduce over NVLink"
        eng.close()
        return out

    def sync_check() -> dict:
        """Short replay on a fresh pipeline without the holdout: the global statistics / model after 3 steps must be
        the same on every rank and equal an NCCL all-reduce of what each rank contributed."""
        if world == 1:
            return None
        F.set_impl("auto")
        params = JobParameters(parallelism=world, jobName="sync_check", testSetSize=0)
        params.extra["sparseDirect"] = "true" if sparse_direct else "false"
        eng = Engine(params, rank, world, dev, impl="fused")
        eng.handle_request({"id": 1, "request": "Create", "learner": {"name": cfg["model"], "hyperParameters": cfg["hyper"]},
                            "preProcessors": cfg["pre"], "trainingConfiguration": tc})
        p = None
        try:
            if cfg["model"] == "ORR" and not sparse:
                n = d + 3
                ref = torch.zeros(n, n, device=dev, dtype=torch.float64)
                for t in range(3):
                    Z = dev_pool[t % n_rot].clone()
                    X = (Z[:, :n] if Z.dtype == torch.float16 else
                         torch.cat([Z[:, : d + 2], torch.zeros(Z.shape[0], 1, device=dev)], 1)).double()
                    ref += X.t() @ X
                    eng.train_batch(Z, d)
                eng.flush()
                dist.all_reduce(ref)                                   # NCCL: sum of the per-rank contributions
                p = eng.pipelines[1]
                lrn = p.learner
                lrn._snapshot()                                        # collective pull of the global statistics
                got = lrn._work[:n, :n].double()
                i = torch.arange(n, device=dev)
                keep = (i[None, :] // 32 * 32 + 31) >= (i[:, None] // 32 * 32)
                err = float(((got - ref).abs() * keep).max() / ref.abs().max())
                g0 = got.clone()
                dist.broadcast(g0, 0)
                spread = torch.tensor([float(((got - g0).abs() * keep).max())], device=dev)
                dist.all_reduce(spread, op=dist.ReduceOp.MAX)
                out = {"what": "global Gram statistics after 3 steps vs NCCL all-reduce of the per-rank Gram matrices (fp64)",
                       "max_rel_err": err, "max_spread_across_ranks": float(spread.item()),
                       "tol": 1e-4, "tol_note": "fp16 products are exact; the fp32 accumulation in TMEM truncates"}
                out["ok"] = bool(err <= out["tol"] and out["max_spread_across_ranks"] == 0.0)
            elif eng.pipelines[1].protocol == "Synchronous":
                for t in range(3):
                    if sparse:
                        eng.train_sparse(dev_pool[t % n_rot].clone(), B, nnz, d)
                    else:
                        eng.train_batch(dev_pool[t % n_rot].clone(), d)
                eng.flush()
                dev_sync()
                p = eng.pipelines[1]
                mine = p.comm.buf.double().clone() if p.learner.merge_mode == "mean" else p.driver.e.double().clone()
                m0 = mine.clone()
                dist.broadcast(m0, 0)
                spread = torch.tensor([float((mine - m0).abs().max())], device=dev)
                dist.all_reduce(spread, op=dist.ReduceOp.MAX)
                out = {"what": "replicas after 3 synchronous steps are identical on every rank",
                       "max_spread_across_ranks": float(spread.item()), "max_rel_err": None}
                out["ok"] = bool(out["max_spread_across_ranks"] == 0.0)
            else:
                out = {"ok": None, "what": f"{eng.pipelines[1].protocol} replicas are not meant to be identical between "
                       "rendezvous; checked by tests/test_multigpu.py instead"}
            out["comm_status"] = eng.comm_status()
        finally:
            eng.close()
        return out

    own_impl = "nccl" if args.impl == "nccl" else "fused"
    ours = measure(own_impl, args.min_timed_ms, not args.no_e2e, True)
    check = None if (dry or own_impl != "fused") else sync_check()
    base = None
    if own_impl == "fused" and not args.no_baseline and not dry:
        try:
            base = measure("nccl", min(args.min_timed_ms, 150.0), not args.no_e2e, False)
        except Exception as ex:
            if world > 1:
                raise
            base = {"error": f"{type(ex).__name__}: {ex}"}
        F.set_impl("auto")

    if rank == 0:
        ratio = (ours["value"] / base["value"]) if (base and base.get("value")) else None
        e2e = ours["e2e"]
        if e2e and base and base.get("e2e") and base["e2e"].get("value") and e2e.get("value"):
            e2e["vs_baseline"] = e2e["value"] / base["e2e"]["value"]
        ld = F.pad32(d + 2)
        dtype = ("fp16 operands (11-bit significand, the precision of a tf32 product), fp32 accumulate in TMEM"
                 if dev_wire == "f16s" else "tf32 (fp32 storage, fp32 accumulate)")
        out = {
            "metric": "records/sec", "value": ours["value"], "unit": "records/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ours["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": ratio,
            "vs_baseline_note": "value / in-repo PyTorch + NCCL (+cuBLAS) baseline measured in this run; NOT the "
                                "reference (a Flink/Kafka job that cannot run here and publishes no numbers)",
            "dtype": dtype, "data": "synthetic", "impl": "ours" if own_impl == "fused" else "nccl-baseline",
            "windows": ours["windows"], "windows_ms": ours["windows_ms"], "window_ms_min": ours["window_ms_min"],
            "window_ms_max": ours["window_ms_max"],
            "config": {"model": cfg["name"], "learner": cfg["model"], "protocol": ours["protocol"],
                       "global_batch": B * world, "per_gpu_batch": B, "features": d, "seq_len": None,
                       "parallelism": f"dp{world} workers + {world} hub shards",
                       "record_format": {"f16s": "fp16 staged rows [x | 1 | y_hi | y_lo], consumed in place",
                                         "f16": "device: fp32 staged rows; e2e: compact fp16 wire rows expanded on the GPU",
                                         "f32": "fp32 staged rows"}[wire],
                       "l2": (f"inputs rotate over {n_rot} distinct sparse slots = "
                              f"{n_rot * B * (2 * nnz + 1) * 4 / 1e6:.0f} MB > 126 MB L2; the model is meant to stay "
                              "L2 resident") if sparse_direct else
                             (f"dense rows are rebuilt every step into 2 alternating {B * ld * 4 / 1e6:.0f} MB buffers "
                              f"(sparse->dense on the GPU); working set per step > 126 MB L2") if sparse else
                             f"inputs rotate over {n_rot} distinct batches = {n_rot * B * row_bytes / 1e6:.0f} MB > 126 MB L2",
                       "sparse_direct": sparse_direct,
                       "sync_every": ours["sync_every"], "model_bytes": ours["model_bytes"],
                       "sync_fusion": str(tc.get("syncFusion", "hub" if ours["driver"] == "ShardedGramDriver" else "none")),
                       "sync_driver": ours["driver"], "sync_transport": ours["transport"]},
            **({"dry_run": "CPU path of the script, not a measurement"} if dry else {}),
            "clocks": ours["clocks"], "numa_node": numa, "e2e": e2e, "gpu_launches": ours["launches"],
            "comm_status": ours["status"], "sync_check": check,
            "baseline": None if base is None else (base if "error" in base else {
                "impl": "in-repo PyTorch + NCCL (+cuBLAS) baseline, same engine and stream, torch ops and "
                        "torch.distributed collectives instead of the fused kernels",
                "value": base["value"], "ms_per_step": base["ms_per_step"], "windows": base["windows"],
                "e2e_value": (base["e2e"] or {}).get("value")}),
        }
        print(json.dumps(out), file=real_stdout, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
