"""Device ML pipeline: ordered preprocessors + one learner + its protocol driver.

A pipeline is what one `Create` request builds (Request.getPreProcessors + getLearner,
PipelineMap.scala:23-30); every worker holds one replica per pipeline
(operators/spoke/SpokeLogic.scala:28-29) and the protocol driver reconciles the replicas.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from ..api.pojos import LearnerPOJO, PreprocessorPOJO, QueryResponse, Request
from ..parallel.comm import PipelineComm
from ..parallel.fused import SyncDriver, make_driver
from ..parallel.host_protocols import cfg
from .learners import GpuLearner
from .learners2 import make_gpu_learner
from .preprocessors import GpuPreprocessor, make_gpu_preprocessor


def effective_protocol(request: Request, world: int, distributed_kmeans: bool) -> str:
    """Protocol rewrites of FlinkSpoke.scala:202-216 / FlinkHub.scala:185-190."""
    tc = request.trainingConfiguration or {}
    proto = tc.get("protocol")
    from ..parallel.fused import DRIVERS
    if proto not in DRIVERS:
        proto = "Asynchronous"                       # MLNodeGenerator.scala:34-43 fallback
    name = request.learner.name
    if name == "HT" or (name == "K-means" and not distributed_kmeans):
        return "SingleLearner"
    if world == 1:
        return "CentralizedTraining"
    return proto


class GpuPipeline:
    def __init__(self, request: Request, device, rank: int, world: int, *, impl: str = "fused", group=None,
                 distributed_kmeans: bool = True, max_msg_params: int = 2000, emulate: int = 1,
                 test_mode: bool = False):
        self.request = request
        self.device = torch.device(device)
        self.rank, self.world, self.impl, self.group, self.emulate = rank, world, impl, group, emulate
        self.max_msg_params = max_msg_params
        self.protocol = effective_protocol(request, world, distributed_kmeans)
        self.preprocessors: List[GpuPreprocessor] = [
            make_gpu_preprocessor(p.name, p.hyperParameters, self.device) for p in (request.preProcessors or [])]
        hyper = dict(request.learner.hyperParameters or {})
        tc = request.trainingConfiguration or {}
        if "trainingMode" in tc and "trainingMode" not in hyper:
            hyper["trainingMode"] = tc["trainingMode"]
        self.learner: GpuLearner = make_gpu_learner(request.learner.name, hyper, self.device)
        self.comm: Optional[PipelineComm] = None
        self.driver: Optional[SyncDriver] = None
        self.d_in: Optional[int] = None
        self.batches = 0
        self.curve: List[tuple] = []
        self._win_loss = 0.0
        self._win_n = 0
        self._inplace = any(p.name != "PolynomialFeatures" for p in self.preprocessors[:1])
        self._fused = None
        self.test_mode = test_mode
        self._curve_every = int(cfg(tc, ["curveEvery", "learningCurveEvery"], 0))
        self._curve_loss, self._curve_n = 0.0, 0

    @property
    def built(self) -> bool:
        return self.comm is not None

    def build(self, d_in: int) -> None:
        """Allocate the symmetric region and bind everything.  COLLECTIVE when world > 1: every rank
        must build its pipelines in the same order."""
        self.d_in = d_in
        d = d_in
        for p in self.preprocessors:
            p.bind(d)
            d = p.out_dim(d)
        lrn = self.learner
        n = lrn.n_flat(d)
        single = self.protocol == "SingleLearner"
        world = 1 if single else self.world
        # HubParallelism (FlinkSpoke.scala:181-190, FlinkHub.scala:176-177): how many parameter-server shards the model
        # is cut into.  The reference defaults to 1; here every GPU is a shard unless the request says otherwise.
        n_hubs = int(cfg(self.request.trainingConfiguration, ["HubParallelism", "hubParallelism"], 0) or 0)
        self.n_hubs = n_hubs if 0 < n_hubs < world else world
        from ..parallel.fused import ShardedGramDriver, wants_scatter, wants_sharded_hub
        scatter_delta = None
        if (wants_sharded_hub(self.protocol, lrn.merge_mode, self.request.trainingConfiguration, world, self.impl,
                              self.device.type) and hasattr(lrn, "use_sharded_hub")):
            # hub sharded over the GPUs, push inside the producing kernel (ShardedGramDriver): buf = this worker's copy of
            # the global statistics, hub = k slot matrices (every worker's running sum of the rows this GPU owns); the
            # worker's own running sum is private to the driver
            ld = int(round(n ** 0.5))                    # n_flat of these learners is ld * ld
            nb, nh = ShardedGramDriver.region_floats(ld, world)
            self.comm = PipelineComm(nb, self.device, self.rank, world, self.group, emulate=1, with_hub=True,
                                     hub_floats=nh, n_hubs=self.n_hubs)
            self.driver = ShardedGramDriver(self.comm, self.comm.buf, lrn.merge_mode, self.request.trainingConfiguration,
                                            self.impl, self.group, self.max_msg_params, ld=ld, n=d + 3)
            lrn.bind(d, self.driver.acc, self.comm.buf)
            self.driver.counters.protocol = self.protocol
            lrn.use_sharded_hub(self.driver)
            self._warm_start(None)
            return
        if wants_scatter(self.protocol, lrn.merge_mode, self.request.trainingConfiguration, world, self.impl) \
                and lrn.needs_e() and n % 4 == 0:
            # reduce-scatter moved into the producer: the learner keeps a private delta, the symmetric region holds
            # three rotating copies of it (parallel/fused.py::ScatterSynchronousDriver)
            self.comm = PipelineComm(3 * n, self.device, self.rank, world, self.group, emulate=1, with_hub=False,
                                     n_hubs=self.n_hubs)
            scatter_delta = torch.zeros(n, device=self.device)
            e = torch.zeros(n, device=self.device)
            lrn.bind(d, scatter_delta, e)
        else:
            self.comm = PipelineComm(n, self.device, self.rank if not single else 0, world, self.group,
                                     emulate=1, with_hub=True, n_hubs=self.n_hubs)
            e = torch.zeros(self.comm.n, device=self.device) if lrn.needs_e() else None
            lrn.bind(d, self.comm.buf, e)
        self.driver = make_driver(self.protocol, self.comm, e, lrn.merge_mode,
                                  self.request.trainingConfiguration, self.impl, self.group, self.max_msg_params,
                                  pipelined_ok=hasattr(lrn, "use_delta_buffers"), scatter_delta=scatter_delta)
        if getattr(self.driver, "gemm_fused", False):
            if hasattr(lrn, "use_scatter_gemm"):
                lrn.use_scatter_gemm(self.driver)
            else:
                self.driver.gemm_fused = False          # learners without a scatter GEMM keep the push kernel
        if hasattr(self.driver, "on_switch"):
            lrn.use_delta_buffers(self.driver.bufs, self.driver.drain)
            self.driver.on_switch = lrn.switch_delta
        self._warm_start(e)

    def _warm_start(self, e) -> None:
        lrn = self.learner
        if self.test_mode and lrn.curve_ring is None and not self._curve_every:
            lrn.enable_curve()           # default test mode: per-mini-batch learning curve kept on the device
        init = self.request.learner.parameters
        if init and init.get("a") is not None:       # warm start from a serialized model
            o = lrn.to_oracle()
            o.from_parameters(init)
            lrn.from_oracle(o)
            if e is not None and lrn.merge_mode == "mean":
                e.copy_(self.comm.buf)

    # ------------------------------------------------------------------ data path
    def preprocess(self, Z: torch.Tensor, fit: bool, protect_input: bool) -> torch.Tensor:
        if self.preprocessors and self._inplace and protect_input:
            Z = Z.clone()
        for p in self.preprocessors:
            Z = p.apply(Z, fit)
        return Z

    def train(self, Z: torch.Tensor, protect_input: bool = False, **kw) -> None:
        if not self.built:
            self.build(self._infer_dim(Z))
        due = self.driver._batches % self.driver.sync_every == self.driver.sync_every - 1
        self.device_train(Z, protect_input, due, **kw)
        self.host_tick()

    def f16_capable(self) -> bool:
        """Can train straight from fp16 staged rows (no expansion to fp32): no preprocessors, a learner with an fp16
        path."""
        return (not self.preprocessors and self.protocol != "SingleLearner"
                and getattr(self.learner, "supports_f16", lambda: False)())

    def sparse_capable(self) -> bool:
        """Can train straight from sparse rows (Engine.train_sparse with `sparseDirect`)."""
        return (not self.preprocessors and self.protocol != "SingleLearner"
                and getattr(self.learner, "supports_sparse", lambda: False)())

    def device_train(self, Z: torch.Tensor, protect_input: bool, sync_due: bool, skip_fit: bool = False,
                     nnz: Optional[int] = None, **kw) -> None:
        """Device work of one step (graph-capturable when `graph_safe()`).  `skip_fit`: the learner
        update was already done by a batched multi-pipeline launch.  `nnz`: Z holds sparse rows."""
        if nnz is not None:
            self.learner.fit_sparse(Z, nnz)
        elif not skip_fit and self._fusable():
            # PolynomialFeatures -> StandardScaler -> mini-batch linear learner: two passes over the raw
            # rows, the expanded features only ever exist in registers (csrc/fused_pipeline.cu)
            from ..ops import functional as F
            pf, sc, l = self.preprocessors[0], self.preprocessors[1], self.learner
            F.fused_poly_scaler_linear(Z, pf.d_in, pf.factors, pf.d_out, sc.state, sc.scratch, l.w, l.stats, l.scratch,
                                       l.kind, l.variant, l.C, l.eps, l.lam, l.mean_update)
        elif not skip_fit:
            Zp = self.preprocess(Z, True, protect_input)
            self.learner.fit_batch(Zp, **kw)
        lrn = self.learner
        if lrn.curve_ring is not None:
            if not lrn._curve_sampled:
                lrn.sample_curve()
            lrn._curve_sampled = False
        if sync_due and self.driver.step():
            self.learner.after_sync()

    def _fusable(self) -> bool:
        if self._fused is None:
            from .learners import GpuLinear
            pp = self.preprocessors
            self._fused = bool(
                len(pp) == 2 and pp[0].name == "PolynomialFeatures" and pp[1].name == "StandardScaler"
                and isinstance(self.learner, GpuLinear) and self.learner.mode != "exact"
                and pp[0].degree <= 4 and pp[0].d_out <= 512
                and str(cfg(self.request.trainingConfiguration, ["fusedPipeline"], "true")).lower() in ("1", "true", "yes"))
        return self._fused

    def host_tick(self) -> None:
        """Host bookkeeping of one step; runs after the device work was enqueued (or replayed)."""
        self.batches += 1
        self.driver.host_tick()
        self.learner.host_tick()
        if self._curve_every and self.batches % self._curve_every == 0:
            # learning-curve sample (loss of the window, #fitted) as the hub records per mini-batch
            # (FlinkHub.scala:101-107); reading the counters synchronises, hence opt-in (`curveEvery`)
            c = self.learner.counters()
            n = c["fitted"] - self._curve_n
            self.curve.append(((c["loss_sum"] - self._curve_loss) / n if n > 0 else 0.0, c["fitted"]))
            self._curve_loss, self._curve_n = c["loss_sum"], c["fitted"]

    def learning_curve(self):
        """Per-mini-batch curve from the device ring (test mode), else the host-side samples (queries, `curveEvery`)."""
        if self.built and self.learner.curve_ring is not None:
            return self.learner.learning_curve()
        return list(self.curve)

    def graph_safe(self) -> bool:
        return (self.built and self.protocol != "SingleLearner" and self.driver.graph_safe
                and self.driver.sync_every == 1 and self.learner.graph_ready())

    def _infer_dim(self, Z: torch.Tensor) -> int:
        d = cfg(self.request.trainingConfiguration, ["nFeatures", "dim"], None)
        if d is None:
            d = (self.request.learner.hyperParameters or {}).get("nFeatures")
        if d is None:
            raise ValueError("feature dimension unknown: pass nFeatures or use Engine.train_batch(Z, d)")
        return int(d)

    def forecast(self, Z: torch.Tensor) -> torch.Tensor:
        return self.learner.predict(self.preprocess(Z, False, True))

    def score(self, ring: torch.Tensor) -> float:
        if not self.built or ring.shape[0] == 0:
            return 0.0
        return self.learner.score(self.preprocess(ring, False, True))

    # ------------------------------------------------------------------ control path
    def local_response(self, response_id: int, ring: torch.Tensor) -> QueryResponse:
        lrn = self.learner
        pps = [PreprocessorPOJO(p.name, dict(p.hyper), p.to_parameters() if self.built else None, None)
               for p in self.preprocessors]
        if not self.built:
            return QueryResponse(response_id, None, self.request.id, pps,
                                 LearnerPOJO(lrn.name, dict(lrn.hyper), None, None), self.protocol, 0, 0.0, 0.0, 0.0)
        c = lrn.counters()
        win_n = c["fitted"] - self._win_n
        loss = (c["loss_sum"] - self._win_loss) / win_n if win_n > 0 else 0.0
        self._win_loss, self._win_n = c["loss_sum"], c["fitted"]
        self.curve.append((loss, c["fitted"]))     # learning curve = (windowed loss, #fitted), FlinkHub.scala:101-107
        pojo = LearnerPOJO(lrn.name, dict(lrn.hyper), lrn.to_parameters(), lrn.data_structure())
        return QueryResponse(response_id, None, self.request.id, pps, pojo, self.protocol, c["fitted"], loss,
                             c["loss_sum"], self.score(ring))

    def state_dict(self) -> dict:
        return {"request": self.request.to_dict(), "d_in": self.d_in, "batches": self.batches,
                "learner": self.learner.state_dict() if self.built else None,
                "pre": [p.state_dict() for p in self.preprocessors] if self.built else None,
                "epoch": self.comm.epoch if self.built else 0, "clock": self.comm.clock if self.built else 0}

    def load_state_dict(self, sd: dict) -> None:
        if sd["learner"] is None:
            return
        if not self.built:
            self.build(sd["d_in"])
        self.learner.load_state_dict(sd["learner"])
        for p, s in zip(self.preprocessors, sd["pre"]):
            p.load_state_dict(s)
        self.batches = sd["batches"]

    def close(self) -> None:
        if self.comm is not None:
            self.comm.close()
            self.comm = None
