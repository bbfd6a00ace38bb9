"""GPU engine: one process per GPU = one worker ("spoke") and one hub shard.

This is the B200-native counterpart of the whole Flink job (Job.scala:35-168,
job/FlinkLearning.scala:33-150).  The public surface a reference user needs is unchanged -
JSON `Create/Update/Query/Delete` requests in, `QueryResponse` / `Prediction` /
`JobStatistics` JSON out - but the data plane is batch oriented: staged record rows come from a
pinned host ring (stand-in for the Kafka trainingData / forecastingData topics), go to the GPU
with one async copy, and every pipeline consumes them with hand-written kernels; replicas are
reconciled by the fused protocol kernels over NVLink (parallel/fused.py, csrc/sync.cu).

SPMD contract: every rank calls `handle_request` with the same requests in the same order and
`train_batch` the same number of times between protocol barriers (the launcher / CLI guarantees
this; requests are broadcast exactly like PipelineMap.scala:54-57 does).
"""
from __future__ import annotations

import contextlib
import os
import time
from typing import Any, Dict, List, Optional, Union

import torch

from .api.bucketing import bucket_response
from .api.pojos import DataInstance, JobStatistics, Prediction, Request, Statistics
from .api.registry import PipelineMap
from .models.pipeline import GpuPipeline
from .ops import functional as F
from .ops import _lib
from .parallel.fused import device_barrier, join_count, query_reduce, split_count
from .stream.parsers import parse_request
from .utils.config import RECORD_BUFFER_CAP, REQUEST_BUFFER_CAP, JobParameters


class Engine:
    def __init__(self, params: Optional[JobParameters] = None, rank: int = 0, world: int = 1,
                 device: Optional[Union[str, torch.device]] = None, impl: str = "fused", group=None):
        self.params = params or JobParameters(parallelism=world)
        self.rank, self.world, self.impl, self.group = rank, world, impl, group
        if device is None:
            device = f"cuda:{torch.cuda.current_device()}" if torch.cuda.is_available() else "cpu"
        self.device = torch.device(device)
        if self.device.type == "cuda" and self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.pipeline_map = PipelineMap(world, self.params.distributedKMeans)
        self.pipelines: Dict[int, GpuPipeline] = {}
        # Creates that arrive before every expected worker is present are parked and replayed once the observed
        # parallelism matches (FlinkSpoke.scala:145-156, requestBuffer cap SpokeLogic.scala:35); `observe_parallelism`
        # is the launcher's hook (torch.distributed gives the final world size at start-up, so by default nothing waits)
        self.pending_requests: List[Request] = []
        self.observed_parallelism = world
        self.expected_parallelism = int(self.params.extra.get("expectedParallelism", world))
        # training records that arrive before the first pipeline exists wait here (recordBuffer, cap 100 000 records,
        # SpokeLogic.scala:32) and are trained on - oldest first - as soon as one does (FlinkSpoke.scala:69-80)
        self._record_buffer: List[torch.Tensor] = []
        self._buffered_rows = 0
        self.dropped_records = 0
        self.holdout: Optional[torch.Tensor] = None   # [cap, ld] staged rows
        self.count = 0                                # training records seen by this worker
        self.d: Optional[int] = None
        self.responses: List[str] = []
        self.predictions: List[str] = []
        self.performance: List[str] = []
        self.t_first: Optional[float] = None
        self.t_last: Optional[float] = None
        self._copy_stream = None
        self._stage: List[Optional[torch.Tensor]] = [None, None]
        self._stage_evt: List[Any] = [None, None]
        self._stage_i = 0
        self._wire_dense: List[Optional[torch.Tensor]] = [None, None]
        self._loss_host = None
        self.last_copy_event = None
        self.holdout_enabled = True
        # CUDA graphs of the steady-state step, one per batch buffer
        self.use_graphs = str(self.params.extra.get("graphs", "true")).lower() in ("1", "true", "yes")
        self.graph_after = 2
        self._graphs: Dict[Any, Any] = {}
        self._graph_launches: Dict[Any, int] = {}
        self._seen: Dict[Any, int] = {}
        self._count_dev: Optional[torch.Tensor] = None
        self._sparse_key = None
        self._hp_stream = None
        # opt-in: measured no gain for the ORR refresh overlap, so the default stays the caller's stream
        self.use_priority_stream = str(self.params.extra.get("priorityStream", "false")).lower() in ("1", "true", "yes")
        self._multi_key = None
        self._multi_desc = None
        # `pipelinedHoldout`: the holdout exchange of batch i + 1 runs on its own stream behind the training kernels of
        # batch i (it only touches batch i + 1 and the ring), so it leaves the critical path of the step.  The caller's
        # stream is then ordered after the engine's work only at the engine's own API boundaries (requests, queries,
        # flush, forecasts, ...), not after every step: a batch tensor handed to train_batch must not be rewritten
        # before the next of those, or before it is handed in again.  Off by default for that reason.
        self.pipelined_holdout = str(self.params.extra.get("pipelinedHoldout", "false")).lower() in ("1", "true", "yes")
        self._hold_stream = None
        self._train_stream = None
        self._pipe_evt: Dict[Any, Any] = {}       # per batch buffer: (inputs ready, holdout done, training done)
        self._user_stream = None
        # steady-state fast path of `_train_rows_now`: batch buffer -> (signature, graph, launches, pipelines to tick);
        # `_epoch` moves whenever a request is processed, which retires every entry
        self._fast: Dict[int, Any] = {}
        self._epoch = 0
        self._nvtx = os.environ.get("OMLDM_NVTX", "0") == "1"     # NVTX ranges around every step (2 us per step)
        self._unsettled = None                    # last training event the caller's stream has not waited for yet
        # sparse streams: train straight from the sparse rows when every pipeline can (csrc/sparse_mcpa.cu)
        # instead of rebuilding the dense batch each step; opt-in until measured on hardware
        self.sparse_direct = str(self.params.extra.get("sparseDirect", "false")).lower() in ("1", "true", "yes")
        self._sparse_nnz: Optional[int] = None
        self._srows: List[Optional[torch.Tensor]] = [None, None]
        self._srows_i = 0
        self._expand: Dict[Any, torch.Tensor] = {}
        self.forward_peer = str(self.params.extra.get("forwardTransport", "peer")).lower() != "nccl"
        self._fwd_comm = None            # SingleLearner forwarding channel (created on first use, collective)
        self._fwd_n = 0
        self._fwd_seq = 0

    # ============================================================================ control plane
    def _settle(self) -> None:
        """Order the caller's stream after everything the engine has in flight (pipelined holdout mode)."""
        if self._unsettled is not None:
            torch.cuda.current_stream(self.device).wait_event(self._unsettled)
            self._unsettled = None

    def handle_request(self, req: Union[str, bytes, dict, Request]) -> List[str]:
        r = parse_request(req)
        if r is None:
            return []
        self._settle()
        self._epoch += 1                  # pipelines may change: every captured step is looked at again
        self._fast.clear()
        out: List[str] = []
        for cm in self.pipeline_map.flat_map(r)[:1]:   # every worker gets the same message; we are one worker
            out += self._process_control(cm.request)
        return out

    def _process_control(self, r: Request) -> List[str]:
        kind = r.request
        self._epoch += 1
        self._fast.clear()
        if kind in ("Create", "Delete"):
            self._graphs.clear()      # captured steps reference the old set of pipelines
            self._graph_launches.clear()
            self._seen.clear()
        if kind == "Create":
            if self.observed_parallelism != self.expected_parallelism:
                if len(self.pending_requests) < REQUEST_BUFFER_CAP:
                    self.pending_requests.append(r)     # parked until every expected worker is present
                return []
            self.pipelines[r.id] = GpuPipeline(
                r, self.device, self.rank, self.world, impl=self.impl, group=self.group,
                distributed_kmeans=self.params.distributedKMeans, max_msg_params=self.params.maxMsgParams,
                test_mode=bool(self.params.test))
            d = (r.trainingConfiguration or {}).get("nFeatures") or (r.learner.hyperParameters or {}).get("nFeatures")
            if d is not None:
                self._build(self.pipelines[r.id], int(d))
            elif self.d is not None:
                self._build(self.pipelines[r.id], self.d)
            return []
        if kind == "Update":
            return []                                   # accepted, routed, no-op (FlinkSpoke.scala:158)
        if kind == "Delete":
            p = self.pipelines.pop(r.id, None)
            if p is not None:
                self._retire(p)
            return []
        if kind == "Query":
            return self.query(r.id, r.requestId)
        return []

    def _retire(self, p: GpuPipeline) -> None:
        """Collective end of a pipeline.  Peers push into this rank's hub shard without a rendezvous under the
        asynchronous protocols, so the symmetric region may only be freed once every rank has stopped using it:
        settle the protocol, device barrier, wait for the stream, then release."""
        if p.built and p.protocol != "SingleLearner":
            p.driver.flush()
            device_barrier(p.comm)
            if self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
            elif self.world > 1:
                import torch.distributed as dist
                if dist.is_available() and dist.is_initialized():
                    dist.barrier(group=self.group)
        p.close()

    def observe_parallelism(self, n: int) -> List[str]:
        """The launcher reports how many workers are present (FlinkSpoke.checkParallelism, :345-348).  Once the
        number matches the expected parallelism the parked Create requests are replayed in arrival order."""
        self.observed_parallelism = int(n)
        out: List[str] = []
        if self.observed_parallelism == self.expected_parallelism and self.pending_requests:
            parked, self.pending_requests = self.pending_requests, []
            for r in parked:
                out += self._process_control(r)
        return out

    def _build(self, p: GpuPipeline, d: int) -> None:
        if not p.built and not (p.protocol == "SingleLearner" and self.rank != 0):
            p.build(d)

    def query(self, pipeline_id: int, response_id: int) -> List[str]:
        self._settle()
        p = self.pipelines.get(pipeline_id)
        if p is None:
            return []
        single = p.protocol == "SingleLearner"
        if single and self.rank != 0:
            return []
        if p.built and not single:
            p.driver.flush()     # collective: pipelined / monitored protocols share their last delta first
        ring = self._holdout_rows()
        resp = p.local_response(response_id, ring)
        if self.world > 1 and not single and p.built:
            lo, mid, hi = split_count(resp.dataFitted or 0)      # exact SUM of the counts (fp32 words, 20-bit chunks)
            vals = torch.tensor([lo, float(resp.loss or 0.0), float(resp.cumulativeLoss or 0.0),
                                 float(resp.score or 0.0), mid, hi], dtype=torch.float32, device=self.device)
            tot = query_reduce(p.comm, vals, self.impl, self.group).tolist()
            resp.dataFitted = join_count(tot[0], tot[4], tot[5])      # SUM (ResponseConstructor.scala:43)
            resp.loss, resp.cumulativeLoss, resp.score = (tot[1] / self.world, tot[2] / self.world,
                                                          tot[3] / self.world)     # MEAN (:44-52)
        if self.rank != 0:
            return []
        frags = [str(f) for f in bucket_response(resp)]
        self.responses += frags
        return frags

    # ============================================================================ data plane
    def _ensure_holdout(self, ld: int, dtype=torch.float32) -> None:
        if self.holdout is None:
            self.holdout = torch.zeros(self.params.testSetSize, ld, device=self.device, dtype=dtype)
        elif self.holdout.dtype != dtype:
            raise ValueError("one stream, one row format: fp32 and fp16 staged rows cannot be mixed in one engine")

    def _holdout_rows(self) -> torch.Tensor:
        if self.holdout is None:
            return torch.zeros(0, 1, device=self.device)
        n = min(F.holdouts_before(self.count), self.holdout.shape[0])
        if self._sparse_nnz is not None:      # the ring holds sparse rows: score on their dense form
            return F.sparse_rows_to_dense(self.holdout[:n], self._sparse_nnz, self.d)
        if self.holdout.dtype == torch.float16:     # fp16 staged rows: score on their fp32 form
            return F.expand_rows_f16s(self.holdout[:n].contiguous(), self.d)
        return self.holdout[:n]

    def train_batch(self, Z: torch.Tensor, d: Optional[int] = None, **kw) -> None:
        """Train every pipeline on one batch of staged rows [B, ld] already on `self.device`.
        The batch is consumed (the holdout swap and in-place scalers modify it).

        Launch-bound steady state: once a batch BUFFER (address + shape) has been seen
        `graph_after` times and every pipeline is graph safe, the device work of the step is captured
        into a CUDA graph and replayed; step-varying scalars (record count, protocol epochs) live in
        device memory so the captured launches stay valid."""
        if d is None:
            d = self.d
        if d is None:
            raise ValueError("feature dimension not known; pass d on the first batch")
        if Z.dim() != 2 or Z.dtype not in (torch.float32, torch.float16) or Z.stride(1) != 1 or Z.stride(0) != Z.shape[1]:
            raise ValueError("train_batch expects a contiguous 2-D tensor of staged rows (fp32, or fp16 'f16s' rows)")
        if Z.dtype == torch.float16:
            if Z.shape[1] != F.f16s_pitch(d):
                raise ValueError(f"fp16 staged rows for d = {d} need a pitch of {F.f16s_pitch(d)} halves "
                                 f"([x | 1 | y_hi | y_lo | pad], omldm_b200.ops.functional.pack_f16s); got {Z.shape[1]}")
        elif Z.shape[1] % 32 or Z.shape[1] < d + 2:
            raise ValueError(f"staged rows need a pitch that is a multiple of 32 floats and >= d + 2 = {d + 2}; "
                             f"got {Z.shape[1]} (use omldm_b200.ops.functional.stage_rows or stream.ring.Topic)")
        if Z.device.type != self.device.type or (Z.device.index is not None and self.device.index is not None
                                                 and Z.device.index != self.device.index):
            raise ValueError(f"batch is on {Z.device}, engine on {self.device}; use train_batch_host for host batches")
        if self.d is not None and self.d != d:
            raise ValueError(f"this stream has {self.d} features, got d = {d}")
        if self._sparse_nnz is not None:
            raise ValueError("this engine trains from sparse rows (sparseDirect); dense batches cannot be mixed in")
        self._train_rows(Z, d, None, **kw)

    def _train_rows(self, Z: torch.Tensor, d: int, nnz: Optional[int], **kw) -> None:
        """One step on validated rows: staged dense rows, or sparse rows when `nnz` is given."""
        self.d = d
        if not self.pipelines:
            self._buffer_records(Z, nnz)           # no pipeline yet: the records wait (FlinkSpoke.scala:79-80)
            return
        if self._record_buffer:
            pending, self._record_buffer, self._buffered_rows = self._record_buffer, [], 0
            for Zb, nb in pending:                 # drain the buffer first, oldest first (FlinkSpoke.scala:72-78)
                self._train_rows_now(Zb, d, nb, **kw)
        self._train_rows_now(Z, d, nnz, **kw)

    def _buffer_records(self, Z: torch.Tensor, nnz: Optional[int]) -> None:
        """Bounded FIFO of whole batches, RECORD_BUFFER_CAP records in total; when it is full the oldest records
        are dropped (DataSet.append evicts the oldest element)."""
        self._record_buffer.append((Z.clone(), nnz))
        self._buffered_rows += Z.shape[0]
        while self._buffered_rows > RECORD_BUFFER_CAP and self._record_buffer:
            Zo, no = self._record_buffer[0]
            excess = self._buffered_rows - RECORD_BUFFER_CAP
            if Zo.shape[0] <= excess:
                self._record_buffer.pop(0)
                self._buffered_rows -= Zo.shape[0]
                self.dropped_records += Zo.shape[0]
            else:
                self._record_buffer[0] = (Zo[excess:].clone(), no)
                self._buffered_rows -= excess
                self.dropped_records += excess

    def _train_rows_now(self, Z: torch.Tensor, d: int, nnz: Optional[int], **kw) -> None:
        now = time.monotonic()
        if self.t_first is None:
            self.t_first = now
        B, ld = Z.shape
        # ---- steady state: this buffer's step is captured and nothing changed since - replay without re-deriving
        # anything (the host side of a step is what bounds small models: ~25 us per step before this path existed)
        fast = self._fast.get(Z.data_ptr()) if not kw else None
        if fast is not None and self.use_graphs and fast[0] == (B, ld, nnz, Z.dtype, self._epoch):
            g, launches, tickers = fast[1], fast[2], fast[3]
            if isinstance(g, tuple):
                self._replay_pipelined(g, Z)
                torch.cuda.set_stream(self._train_stream)      # cadence work is ordered after this step's kernels
                try:
                    for p in tickers:
                        p.host_tick()
                finally:
                    torch.cuda.set_stream(self._user_stream)
            else:
                if self._nvtx:
                    torch.cuda.nvtx.range_push("omldm.step.replay")
                g.replay()
                if self._nvtx:
                    torch.cuda.nvtx.range_pop()
                for p in tickers:
                    p.host_tick()
            _lib.add_launch_count(launches)
            self.count += B
            self.t_last = now
            return
        for p in self.pipelines.values():
            if p.protocol != "SingleLearner" or self.rank == 0:
                self._build(p, d)
        keyed = [p for p in self.pipelines.values() if p.built and hasattr(p.driver, "graph_key")]
        key = (Z.data_ptr(), B, ld, nnz, Z.dtype) + tuple(p.driver.graph_key() for p in keyed)
        g = None
        if self.use_graphs and not kw and self._graphable(Z):
            g = self._graphs.get(key)
            if g is None and self._seen.get(key, 0) >= self.graph_after:
                try:
                    g = self._capture(key, Z, d, nnz)
                except Exception as e:      # never lose a job over an optimisation: fall back to eager steps
                    import warnings
                    warnings.warn(f"CUDA graph capture failed ({e}); continuing with eager launches")
                    self.use_graphs = False
                    g = None
        cuda = self.device.type == "cuda"
        if cuda and self._nvtx:
            torch.cuda.nvtx.range_push("omldm.step.replay" if g is not None else "omldm.step.eager")
        hp = self._step_stream() if cuda else None
        if hp is not None:
            # the step runs on a high-priority stream so that background work on ordinary streams (the
            # ORR weight refresh, a checkpoint copy) cannot push the hot kernels out of the SMs
            user = torch.cuda.current_stream(self.device)
            hp.wait_stream(user)
            ctx = torch.cuda.stream(hp)
            ctx.__enter__()
        pipelined = isinstance(g, tuple)
        try:
            if pipelined:
                self._replay_pipelined(g, Z)
                _lib.add_launch_count(self._graph_launches[key])
            elif g is not None:
                g.replay()
                _lib.add_launch_count(self._graph_launches[key])
            else:
                self._settle()
                self._seen[key] = self._seen.get(key, 0) + 1
                self._device_step(Z, d, nnz, **kw)
        finally:
            if hp is not None:
                ctx.__exit__(None, None, None)
                user.wait_stream(hp)
        if cuda and self._nvtx:
            torch.cuda.nvtx.range_pop()
        self.count += B
        # cadence work of the learners / drivers (weight refresh, FGM rounds): ordered after this step's kernels, so in
        # pipelined mode it is issued on the training stream
        tickers = [p for p in self.pipelines.values() if p.built and (p.protocol != "SingleLearner" or self.rank == 0)]
        tick_ctx = torch.cuda.stream(self._train_stream) if pipelined else contextlib.nullcontext()
        with tick_ctx:
            for p in tickers:
                p.host_tick()
        if g is not None and not keyed and hp is None:
            # from now on this buffer takes the fast path, until a request changes the set of pipelines (`_epoch`)
            self._fast[Z.data_ptr()] = ((B, ld, nnz, Z.dtype, self._epoch), g, self._graph_launches[key], tickers)
        self.t_last = time.monotonic()

    def _step_stream(self):
        if not self.use_priority_stream:
            return None
        if self._hp_stream is None:
            self._hp_stream = torch.cuda.Stream(device=self.device, priority=-1)
        return self._hp_stream

    def _graphable(self, Z: torch.Tensor) -> bool:
        if not (Z.is_cuda and F.native(Z)) or not self.pipelines:
            return False
        return all(p.graph_safe() for p in self.pipelines.values())

    def _capture(self, key, Z: torch.Tensor, d: int, nnz: Optional[int] = None):
        torch.cuda.synchronize(self.device)
        before = _lib.launch_count()
        if self.pipelined_holdout and self.holdout_enabled and self.params.testSetSize > 0:
            # two graphs: the holdout exchange (replayed on the holdout stream, behind the previous batch's training
            # kernels) and the training step proper
            gh, g = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
            with torch.cuda.graph(gh):
                self._holdout_step(Z, d)
            with torch.cuda.graph(g):
                self._device_step(Z, d, nnz, skip_holdout=True)
            g = (gh, g)
        else:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._device_step(Z, d, nnz)
        self._graph_launches[key] = _lib.launch_count() - before
        _lib.add_launch_count(-self._graph_launches[key])   # capture does not execute anything
        self._graphs[key] = g
        return g

    def _replay_pipelined(self, g, Z: torch.Tensor) -> None:
        """Holdout graph on the holdout stream, training graph on the training stream; see `pipelinedHoldout`.
        (set_stream instead of stream contexts: 1 us against 7 us each on the host.)"""
        gh, gt = g
        if self._hold_stream is None:
            self._hold_stream = torch.cuda.Stream(device=self.device)
            self._train_stream = torch.cuda.Stream(device=self.device)
        ev = self._pipe_evt.get(Z.data_ptr())
        if ev is None:
            ev = self._pipe_evt[Z.data_ptr()] = (torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event(), [False])
        e_in, e_hold, e_train, used = ev
        user = self._user_stream = torch.cuda.current_stream(self.device)
        sh, st = self._hold_stream, self._train_stream
        if used[0]:
            sh.wait_event(e_train)            # the previous training step on THIS buffer has finished reading it
        e_in.record(user)                     # whatever the caller enqueued before this call (its input, a request)
        sh.wait_event(e_in)
        st.wait_event(e_in)
        try:
            torch.cuda.set_stream(sh)
            gh.replay()
            e_hold.record(sh)
            st.wait_event(e_hold)
            torch.cuda.set_stream(st)
            gt.replay()
            e_train.record(st)
        finally:
            torch.cuda.set_stream(user)
        used[0] = True
        self._unsettled = e_train

    def _holdout_step(self, Z: torch.Tensor, d: int) -> None:
        """The 2-in-10 holdout rule on one batch, in place (FlinkSpoke.scala:92-107)."""
        if not (self.holdout_enabled and self.params.testSetSize > 0):
            return
        ld = Z.shape[1]
        half = Z.dtype == torch.float16
        self._ensure_holdout(ld, Z.dtype)
        # fp16 staged rows move as raw bytes: the holdout rule is a pure row move, and an all-zero row is empty in
        # both layouts, so the same kernel serves both with the rows viewed as fp32 words
        Zv, ring = (Z.view(torch.float32), self.holdout.view(torch.float32)) if half else (Z, self.holdout)
        if F.native(Z):                     # (true for CUDA batches; also under the CPU kernel emulation)
            if self._count_dev is None:
                self._count_dev = torch.tensor([self.count, 0], dtype=torch.int64, device=self.device)
            # 2-in-10 holdout, evict-then-train; per-record learners need the sequential positions, mini-batch
            # learners only the set (one exchange per ring slot)
            ordered = any(getattr(p.learner, "mode", "") == "exact" for p in self.pipelines.values())
            F.holdout_swap_dev(Zv, ring, d, self._count_dev, ordered=ordered)
        else:
            F.holdout_swap(Zv, ring, d, self.count)

    def _device_step(self, Z: torch.Tensor, d: int, nnz: Optional[int] = None, skip_holdout: bool = False, **kw) -> None:
        """All device work of one training step, in stream order (`skip_holdout`: the holdout rule has already been
        applied to this batch on the holdout stream)."""
        B, ld = Z.shape
        if nnz is not None:
            kw = dict(kw, nnz=nnz)
        for p in self.pipelines.values():
            if p.built and hasattr(p.driver, "begin_step"):
                p.driver.begin_step()          # fork: overlapped all-reduce of the previous step's delta
        half = Z.dtype == torch.float16
        if not skip_holdout:
            self._holdout_step(Z, d)
        Z32 = None
        if half and not all(p.f16_capable() for p in self.pipelines.values() if p.built or p.protocol != "SingleLearner"):
            # some pipeline has no fp16 path: expand once per step (csrc/rows_f16.cu)
            kx = (B, d)
            if self._expand.get(kx) is None:
                self._expand = {kx: torch.empty(B, F.pad32(d + 2), device=self.device)}
            Z32 = F.expand_rows_f16s(Z, d, out=self._expand[kx])
        n_pipes = len(self.pipelines)
        batched = self._batched_exact_fit(Z32 if half else Z, d) if (nnz is None and (not half or Z32 is not None)) else ()
        for p in self.pipelines.values():
            due = p.built and (p.driver._batches % p.driver.sync_every == p.driver.sync_every - 1)
            Zp = Z if (not half or p.f16_capable()) else Z32
            if p in batched:
                p.device_train(Zp, False, due, skip_fit=True)
                continue
            if p.protocol == "SingleLearner":
                Zs = self._forward_to_rank0(Zp)
                if self.rank == 0:
                    p.device_train(Zs, True, due, **kw)
                continue
            p.device_train(Zp, n_pipes > 1 and not (half and Zp is Z), due, **kw)

    def _batched_exact_fit(self, Z: torch.Tensor, d: int):
        """Multi-tenant fast path: all exact-mode linear pipelines without preprocessors advance in ONE
        kernel launch, one warp per pipeline (tiny models are latency bound; the launch is shared)."""
        from .models.learners import GpuLinear
        if not F.native(Z):
            return ()
        group = [p for p in self.pipelines.values()
                 if p.built and p.protocol != "SingleLearner" and isinstance(p.learner, GpuLinear)
                 and p.learner.mode == "exact" and not p.preprocessors]
        if len(group) < 2:
            return ()
        key = tuple(id(p) for p in group)
        if self._multi_key != key:
            self._multi_key = key
            self._multi_desc = F.make_linear_descs(
                [(p.learner.w, p.learner.stats, p.learner.d, p.learner.kind, p.learner.variant, p.learner.C,
                  p.learner.eps, p.learner.lam) for p in group], self.device)
        import contextlib
        with (torch.cuda.nvtx.range("linear_exact_multi") if Z.is_cuda else contextlib.nullcontext()):
            F.linear_exact_multi(Z, d, self._multi_desc, len(group))
        return group

    def _forward_to_rank0(self, Z: torch.Tensor) -> torch.Tensor:
        """SingleLearner: workers forward raw records to the one learner at hub 0."""
        if self.world == 1:
            return Z
        if self.impl == "fused" and Z.is_cuda and F.native(Z) and self.forward_peer:
            return self._forward_fused(Z)
        import torch.distributed as dist
        if self.rank == 0:
            parts = [torch.empty_like(Z) for _ in range(self.world)]
            dist.gather(Z, parts, dst=0, group=self.group)
            return torch.cat(parts, 0)
        dist.gather(Z, None, dst=0, group=self.group)
        return Z

    def _forward_fused(self, Z: torch.Tensor) -> torch.Tensor:
        """Forwarding over NVLink peer memory (csrc/sync.cu::sync_forward_kernel): every rank stores its batch into
        rank 0's symmetric buffer, rank 0 gets a view of the world concatenated batches.  Collective; every rank
        must pass batches of the same shape."""
        import ctypes
        from .parallel.comm import PipelineComm
        B, ld = Z.shape
        n = B * ld
        if self._fwd_comm is None or self._fwd_n != n:
            if self._fwd_comm is not None:
                self._fwd_comm.close()
            self._fwd_comm = PipelineComm(2 * self.world * n, self.device, self.rank, self.world, self.group,
                                          emulate=1, with_hub=False)
            self._fwd_n, self._fwd_seq = n, 0
        self._fwd_seq += 1
        c = self._fwd_comm
        _lib.check(_lib.kernels().omldm_sync_forward(c.struct_ptr(), _lib.ptr(Z), ctypes.c_size_t(n),
                                                     ctypes.c_uint32(self._fwd_seq), ctypes.c_int(0),
                                                     _lib.stream_ptr()), "sync_forward")
        if self.rank != 0:
            return Z
        return c.buf.view(2, self.world * B, ld)[self._fwd_seq & 1]

    # ---- end-to-end path: pinned host slot -> device -> train -> loss back to the host -------------
    def _stage_slot(self, host: torch.Tensor):
        """Async H2D copy of a pinned host slot into one of two alternating device buffers.  The copy waits only for
        the step that last read THAT buffer (two steps ago), so the copy of batch t + 1 overlaps the compute of
        batch t.  Returns (buffer index, device tensor); call `_stage_done(i)` after enqueueing the step."""
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        i = self._stage_i
        self._stage_i ^= 1
        cur = torch.cuda.current_stream(self.device)
        if self._stage[i] is None or self._stage[i].shape != host.shape or self._stage[i].dtype != host.dtype:
            self._stage[i] = torch.empty(host.shape, dtype=host.dtype, device=self.device)
            self._stage_evt[i] = None
            self._copy_stream.wait_stream(cur)
        if self._stage_evt[i] is not None:
            self._copy_stream.wait_event(self._stage_evt[i])
        with torch.cuda.stream(self._copy_stream):
            self._stage[i].copy_(host, non_blocking=True)
            self.last_copy_event = torch.cuda.Event()
            self.last_copy_event.record(self._copy_stream)
        cur.wait_event(self.last_copy_event)
        return i, self._stage[i]

    def _buffer_free(self, t: torch.Tensor) -> None:
        """The current stream may rewrite an engine-owned batch buffer once the last training step that read it has
        finished (pipelined holdout mode; otherwise stream order already says so)."""
        ev = self._pipe_evt.get(t.data_ptr())
        if ev is not None and ev[3][0]:
            torch.cuda.current_stream(self.device).wait_event(ev[2])

    def _stage_done(self, i: int) -> None:
        self._settle()
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._stage_evt[i] = ev

    def train_batch_host(self, Z_host: torch.Tensor, d: int, wire: Optional[str] = None, **kw) -> torch.Tensor:
        """One e2e step from a pinned host slot.  Returns a pinned fp64[4] tensor that will hold
        (loss_sum, aux, fitted, extra) of the first pipeline once the current stream has drained.
        `wire` names the slot's row format ("f32" staged rows, "f16" compact wire rows, "f16s" fp16 staged rows; see
        stream.ring.Topic); it is inferred from dtype and pitch when it is not given."""
        if Z_host.dtype == torch.float16:
            if wire is None:
                compact, staged = Z_host.shape[1] == F.wire_f16_pitch(d), Z_host.shape[1] == F.f16s_pitch(d)
                if compact and staged:
                    raise ValueError(f"for d = {d} the compact and the staged fp16 formats have the same pitch: pass wire=")
                wire = "f16s" if staged else "f16"
            if wire == "f16":
                return self._train_batch_host_f16(Z_host, d, **kw)
        if self.device.type != "cuda":
            self.train_batch(Z_host.clone(), d, **kw)
            return self._read_stats_host()
        i, Zd = self._stage_slot(Z_host)
        self.train_batch(Zd, d, **kw)
        self._stage_done(i)
        return self._read_stats_host()

    def _train_batch_host_f16(self, W_host: torch.Tensor, d: int, **kw) -> torch.Tensor:
        """e2e step from a slot in the compact fp16 wire format (`stream.ring.Topic(wire="f16")`):
        half the H2D bytes, expanded to staged fp32 rows on the GPU (csrc/ingest.cu::stage_rows_f16_kernel)
        into one of two alternating buffers, then the normal step."""
        if W_host.shape[1] != F.wire_f16_pitch(d):
            raise ValueError(f"fp16 wire rows for d={d} must have {F.wire_f16_pitch(d)} halves per record, "
                             f"got {W_host.shape[1]}")
        if self.device.type != "cuda":
            self.train_batch(F.stage_rows_f16(W_host, d), d, **kw)
            return self._read_stats_host()
        i, Wd = self._stage_slot(W_host)
        B, ld = W_host.shape[0], F.pad32(d + 2)
        if self._wire_dense[i] is None or self._wire_dense[i].shape != (B, ld):
            self._wire_dense[i] = torch.empty(B, ld, dtype=torch.float32, device=self.device)
        self._buffer_free(self._wire_dense[i])
        F.stage_rows_f16(Wd, d, out=self._wire_dense[i])
        self.train_batch(self._wire_dense[i], d, **kw)
        self._stage_done(i)
        return self._read_stats_host()

    # ---- sparse streams (BASELINE config 5: "100k-dim sparse-to-dense stream") ---------------------
    def train_sparse(self, slot: torch.Tensor, batch: int, nnz: int, d: int) -> None:
        """One step from a fixed-width sparse slot on the device ([int32 idx | float val | float y],
        see stream.ring.SparseTopic): densify on the GPU (csrc/ingest.cu), then the normal step."""
        from .stream.ring import SparseTopic
        idx, val, tgt = SparseTopic.unpack(slot, batch, nnz)
        if self.sparse_direct and self._sparse_capable(d):
            if self.holdout is not None and self._sparse_nnz is None:
                raise ValueError("this engine already trained on dense batches; sparseDirect cannot be mixed in")
            if self.d is not None and self.d != d:
                raise ValueError(f"this stream has {self.d} features, got d = {d}")
            i = self._srows_i
            self._srows_i ^= 1
            shape = (batch, F.sparse_row_pitch(nnz))
            if self._srows[i] is None or self._srows[i].shape != shape:
                self._srows[i] = torch.empty(shape, dtype=torch.float32, device=self.device)
            self._sparse_nnz = nnz
            self._buffer_free(self._srows[i])
            F.sparse_rows(idx, val, tgt, batch, nnz, out=self._srows[i])
            self._train_rows(self._srows[i], d, nnz)
            return
        key = (batch, nnz, d)
        if self._sparse_key != key:
            self._sparse_key = key
            self._indptr = torch.arange(0, (batch + 1) * nnz, nnz, dtype=torch.int32, device=self.device)
            self._dense = [torch.empty(batch, F.pad32(d + 2), device=self.device) for _ in range(2)]
            self._sqn = torch.empty(batch, device=self.device)
            self._dense_i = 0
        Z = self._dense[self._dense_i]
        self._dense_i ^= 1
        self._buffer_free(Z)
        F.densify(self._indptr, idx, val, tgt, d, out=Z, sq_out=self._sqn)
        self.train_batch(Z, d)

    def _sparse_capable(self, d: int) -> bool:
        """Every pipeline can consume sparse rows (no preprocessors, a learner with a sparse step)."""
        ps = list(self.pipelines.values())
        return bool(ps) and all(p.sparse_capable() for p in ps)

    def train_sparse_host(self, slot_host: torch.Tensor, batch: int, nnz: int, d: int) -> torch.Tensor:
        """e2e sparse step: pinned host slot -> H2D (a few hundred KB instead of the dense batch) ->
        densify -> train; returns the pinned stats tensor like `train_batch_host`."""
        if self.device.type != "cuda":
            self.train_sparse(slot_host.clone(), batch, nnz, d)
            return self._read_stats_host()
        i, slot = self._stage_slot(slot_host)
        self.train_sparse(slot, batch, nnz, d)
        self._stage_done(i)
        return self._read_stats_host()

    def _read_stats_host(self) -> torch.Tensor:
        self._settle()
        if self._loss_host is None:
            self._loss_host = torch.zeros(4, dtype=torch.float64)
            if self.device.type == "cuda":
                self._loss_host = self._loss_host.pin_memory()
        for p in self.pipelines.values():
            if p.built:
                self._loss_host.copy_(p.learner.stats, non_blocking=True)
                break
        return self._loss_host

    def flush(self) -> None:
        """Collective end-of-stream hook: let every protocol driver settle (share the last pipelined
        delta, resolve an FGM round that only some ranks have entered).  Call it on every rank before
        a host-side barrier, otherwise a rank may sit in a device-side rendezvous that its peers will
        never join."""
        self._settle()
        for p in self.pipelines.values():
            if p.built and p.protocol != "SingleLearner":
                p.driver.flush()

    def forecast_batch(self, Z: torch.Tensor, instances: Optional[List[DataInstance]] = None) -> List[str]:
        """Every pipeline predicts every forecasting record with its local replica
        (FlinkSpoke.scala:105, FlinkNetwork.scala:250)."""
        self._settle()
        out: List[str] = []
        for pid, p in self.pipelines.items():
            if not p.built:
                continue
            preds = p.forecast(Z).tolist()
            for i, v in enumerate(preds):
                di = instances[i] if instances else None
                out.append(str(Prediction(pid, di, float(v))))
        self.predictions += out
        return out

    # ============================================================================ termination / stats
    def statistics(self) -> List[Statistics]:
        self._settle()
        out = []
        ring = self._holdout_rows()
        for pid, p in self.pipelines.items():
            if not p.built:
                continue
            c = p.learner.counters()
            k = p.driver.counters
            out.append(Statistics(pipeline=pid, protocol=k.protocol, models=k.models, bytes=k.bytes, blocks=k.blocks,
                                  fitted=c["fitted"], score=p.score(ring),
                                  learningCurve=[x[0] for x in p.learning_curve()],
                                  lcx=[x[1] for x in p.learning_curve()]))
        return out

    def terminate(self) -> Optional[JobStatistics]:
        """Termination probe: device barrier, final holdout evaluation, JobStatistics on rank 0
        (StatisticsOperator.scala:93-142)."""
        self._settle()
        for p in self.pipelines.values():
            if p.built and p.protocol != "SingleLearner":
                p.driver.flush()
                device_barrier(p.comm)
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        stats = self.statistics()
        dur_ms = 1000.0 * ((self.t_last or 0.0) - (self.t_first or 0.0))
        if self.world > 1:
            import torch.distributed as dist
            gathered: List[Any] = [None] * self.world
            dist.all_gather_object(gathered, [s.to_dict() for s in stats], group=self.group)
            if self.rank != 0:
                return None
            merged: Dict[int, Statistics] = {}
            for per_rank in gathered:
                for sd in per_rank:
                    m = merged.setdefault(sd["pipeline"], Statistics(sd["pipeline"], sd["protocol"]))
                    m.fitted += sd["fitted"]
                    m.score += sd["score"] / self.world
                    m.models, m.bytes, m.blocks = max(m.models, sd["models"]), m.bytes + sd["bytes"], \
                        max(m.blocks, sd["blocks"])
                    if not m.learningCurve:
                        m.learningCurve, m.lcx = sd["learningCurve"], sd["lcx"]
            stats = list(merged.values())
        js = JobStatistics(self.params.jobName, self.world, dur_ms, stats)
        if dur_ms > 0:
            js.extra["recordsPerSec"] = sum(s.fitted for s in stats) / (dur_ms / 1000.0)
            js.extra["syncBytesPerSec"] = sum(s.bytes for s in stats) / (dur_ms / 1000.0)
        self.performance.append(str(js))
        return js

    def comm_status(self) -> int:
        st = 0
        for p in self.pipelines.values():
            if p.built and p.comm.device.type == "cuda":
                st |= p.comm.status()
        if self._fwd_comm is not None and self.device.type == "cuda":
            st |= self._fwd_comm.status()
        return st

    def start_watchdog(self, period_s: float = 0.5, stall_after_s: float = 5.0, on_stall=None):
        """Peer liveness (utils/watchdog.py): heartbeat counters in the signal pads of the first distributed pipeline.
        Returns the running `Watchdog`, or None when there is nothing to watch (one rank, no pipeline built yet)."""
        from .utils.watchdog import Watchdog
        for p in self.pipelines.values():
            if p.built and p.comm.k > 1:
                self.watchdog = Watchdog(p.comm, self.rank, period_s, stall_after_s, on_stall).start()
                return self.watchdog
        return None

    def health(self) -> Dict[str, Any]:
        """Timeout bits of the protocol kernels and, if a watchdog runs, the heartbeat picture."""
        wd = getattr(self, "watchdog", None)
        return {"rank": self.rank, "comm_status": self.comm_status(), "watchdog": None if wd is None else wd.snapshot()}

    # ============================================================================ checkpoint
    def state_dict(self) -> dict:
        self._settle()
        return {"rank": self.rank, "world": self.world, "count": self.count, "d": self.d,
                "registry": self.pipeline_map.state_dict(),
                "holdout": None if self.holdout is None else self.holdout.cpu().clone(),
                "sparse_nnz": self._sparse_nnz,
                "pipelines": {pid: p.state_dict() for pid, p in self.pipelines.items()}}

    def load_state_dict(self, sd: dict) -> None:
        self.count, self.d = sd["count"], sd["d"]
        self._sparse_nnz = sd.get("sparse_nnz")
        self._count_dev = None
        self._graphs, self._seen = {}, {}
        self._fast.clear()
        self._epoch += 1
        self.pipeline_map.load_state_dict(sd["registry"])
        if sd["holdout"] is not None:
            self.holdout = sd["holdout"].to(self.device)
        for pid, ps in sd["pipelines"].items():
            req = Request.from_dict(ps["request"])
            if pid not in self.pipelines:
                self._process_control(req)
            self.pipelines[pid].load_state_dict(ps)

    def close(self) -> None:
        self._settle()
        if getattr(self, "watchdog", None) is not None:
            self.watchdog.stop()
            self.watchdog = None
        for p in self.pipelines.values():
            p.close()
        self.pipelines = {}
        if self._fwd_comm is not None:
            self._fwd_comm.close()
            self._fwd_comm = None
