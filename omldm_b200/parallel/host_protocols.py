"""Message-driven worker / parameter-server nodes for the eight synchronisation protocols.

This is the host-side (CPU, fp64) realisation of the out-of-tree `mlAPI.mlworkers.*` /
`mlAPI.mlParameterServers.*` classes that the reference instantiates by protocol name
(utils/generators/MLNodeGenerator.scala:25-36 worker side, :54-65 hub side; unknown names fall
back to Asynchronous :34-43,:63-72).  It runs on any `Network` implementation
(`parallel/network.py`), so the protocol invariants are testable on a CPU-only box; the GPU
engine (`engine.py` + csrc/sync.cu) implements the same semantics with fused kernels.

Protocol semantics are the literature ones: Synchronous = BSP averaging,
Asynchronous = fold-in on arrival + reply to sender, SSP (Ho et al. 2013), EASGD (Zhang et al.
2015), GM (Sharfman et al. 2006, ball safe zone), FGM (Samoladas & Garofalakis 2019).

Model sharding: the flat parameter vector is cut into H contiguous shards, shard h lives on hub
h (`HubParallelism`, FlinkSpoke.scala:184, FlinkHub.scala:176-177).  Messages carry at most
`maxMsgParams` parameters (DefaultJobParameters.scala:6); the pieces are the "blocks" counted in
the statistics (FlinkHub.scala:123).
"""
from __future__ import annotations

import math
from typing import Any, Dict, List, Optional, Tuple

import numpy as np

from ..api.compat import double_parsing, integer_parsing
from ..api.pojos import Prediction, Request
from ..messages import NodeId, NodeType, RemoteCallIdentifier
from ..oracle.pipeline import OraclePipeline

PROTOCOLS = ("CentralizedTraining", "SingleLearner", "Asynchronous", "Synchronous", "SSP", "EASGD", "GM", "FGM")


def cfg(tc: Optional[dict], keys, default):
    if tc:
        for k in keys if isinstance(keys, (list, tuple)) else [keys]:
            if k in tc and tc[k] is not None:
                return tc[k]
    return default


def cfg_int(tc: Optional[dict], keys, default: int) -> int:
    """Typed lookup with the reference's `Parsing.IntegerParsing` rule: absent or malformed -> the default."""
    return integer_parsing({"v": cfg(tc, keys, None)}, "v", default)


def cfg_float(tc: Optional[dict], keys, default: float) -> float:
    return double_parsing({"v": cfg(tc, keys, None)}, "v", default)


def shard_bounds(n: int, H: int) -> List[Tuple[int, int]]:
    per = (n + H - 1) // H if H > 0 else n
    return [(min(n, h * per), min(n, (h + 1) * per)) for h in range(H)]


class ProtocolStatistics:
    def __init__(self, protocol: str):
        self.protocol = protocol
        self.models_shipped = 0
        self.bytes_shipped = 0
        self.num_of_blocks = 0

    def ship(self, n_params: int, blocks: int = 1, models: int = 1) -> None:
        self.models_shipped += models
        self.bytes_shipped += 8 * n_params
        self.num_of_blocks += blocks


# ============================================================================== workers
class WorkerNode:
    """Base of all worker classes (`NodeInstance`, utils/generators/NodeGenerator.scala:8-9)."""
    protocol = "Asynchronous"

    def __init__(self, max_msg_params: int = 2000):
        self.max_msg_params = max_msg_params
        self.node_id: Optional[NodeId] = None
        self.network = None
        self.pipeline: Optional[OraclePipeline] = None
        self.request: Optional[Request] = None
        self.count = 0
        self.blocked = False
        self.e: Optional[np.ndarray] = None     # last global estimate installed locally
        self._pending: Dict[int, Any] = {}
        self._pushes = 0

    # -- `.configureWorker(request)` (MLNodeGenerator.scala:26-35)
    def configure_worker(self, request: Request) -> "WorkerNode":
        self.request = request
        self.pipeline = OraclePipeline(request)
        tc = request.trainingConfiguration
        self.mini_batch = cfg_int(tc, ["miniBatchSize", "mini_batch_size", "batchSize"], 64)
        self.H = cfg_int(tc, ["HubParallelism", "hubParallelism"], 1)
        self.tc = tc or {}
        return self

    def bind(self, node_id: NodeId, network) -> None:
        self.node_id, self.network = node_id, network

    # ------------------------------------------------------------------ data path
    def receive_tuple(self, point) -> None:
        from ..stream.points import ForecastingPoint, TrainingPoint
        if isinstance(point, ForecastingPoint):
            self.forecast(point)
            return
        assert isinstance(point, TrainingPoint)
        if self.e is None:
            # eager init so that e starts at the (possibly random) initial model, not at zero
            xx = self.pipeline._pre(np.asarray(point.numerical, dtype=np.float64), False)
            self.pipeline.learner.ensure_init(len(xx))
            self.e = self.flat.copy()
        self.pipeline.fit(point.numerical, point.target)
        self.count += 1
        if self.count % self.mini_batch == 0:
            self.on_mini_batch()

    def forecast(self, point) -> None:
        pred = self.pipeline.predict(point.numerical)
        self.network.send(self.node_id, None, RemoteCallIdentifier("prediction"),
                          Prediction(self.request.id, point.data_instance, pred))

    def on_mini_batch(self) -> None:
        pass

    # ------------------------------------------------------------------ control path
    def receive_msg(self, source: NodeId, rpc: RemoteCallIdentifier, data: Any) -> None:
        getattr(self, "rpc_" + rpc.op)(source, data)

    def receive_query(self, response_id: int, payload) -> None:
        mean_buf, test = payload
        X, y = test
        resp = self.pipeline.query_response(response_id, self.protocol, (X, y))
        if response_id == -1:
            resp.cumulativeLoss = float(mean_buf)  # FlinkLearning.scala:122 reads it back as buffer size
        self.network.send(self.node_id, None, RemoteCallIdentifier("query_response"), resp)

    # ------------------------------------------------------------------ helpers
    @property
    def flat(self) -> np.ndarray:
        return self.pipeline.learner.get_flat()

    def ready(self) -> bool:
        return self.pipeline.learner.d is not None

    def n_workers(self) -> int:
        return self.network.describe().n_spokes

    def scale(self) -> float:
        return 1.0 if self.pipeline.learner.merge_mode == "sum" else 1.0 / self.n_workers()

    def push(self, op: str, vec: np.ndarray, extra: Optional[dict] = None) -> None:
        """Send shard h of `vec` to hub h in blocks of <= max_msg_params parameters."""
        bounds = shard_bounds(len(vec), self.H)
        for h, (lo, hi) in enumerate(bounds):
            n = hi - lo
            nblocks = max(1, math.ceil(n / self.max_msg_params))
            for b in range(nblocks):
                s = lo + b * self.max_msg_params
                t = min(hi, s + self.max_msg_params)
                payload = {"offset": s - lo, "values": vec[s:t].copy(), "shard_len": n, "total": len(vec),
                           "last": b == nblocks - 1}
                if self._pushes == 0 and self.e is not None and len(self.e) == len(vec):
                    payload["base"] = self.e[s:t].copy()   # lets the hub start from the initial model
                if extra:
                    payload.update(extra)
                self.network.send(self.node_id, NodeId(NodeType.HUB, h), RemoteCallIdentifier(op), payload)
        self._pushes += 1

    def install_shard(self, source: NodeId, data: dict) -> bool:
        """Collect one hub's shard; True once every hub's shard for this round arrived."""
        self._pending[source.node_id] = data
        if len(self._pending) < self.H:
            return False
        total = next(iter(self._pending.values()))["total"]
        vec = np.zeros(total)
        for h, (lo, hi) in enumerate(shard_bounds(total, self.H)):
            vec[lo:hi] = self._pending[h]["values"]
        self._pending = {}
        self.pipeline.learner.set_flat(vec)
        self.e = vec.copy()
        return True

    def ensure_e(self) -> None:
        if self.e is None or len(self.e) != len(self.flat):
            self.e = np.zeros_like(self.flat)


class SingleWorker(WorkerNode):
    """`CentralizedTraining`: forced when parallelism == 1 (FlinkSpoke.scala:213-216)."""
    protocol = "CentralizedTraining"

    def on_mini_batch(self):
        # the lone learner only reports progress so the hub can keep statistics / learning curve
        lrn = self.pipeline.learner
        self.network.send(self.node_id, NodeId(NodeType.HUB, 0), RemoteCallIdentifier("report"),
                          {"fitted": self.mini_batch, "loss": self.pipeline.window_loss / max(1, self.pipeline.window_n),
                           "cum": lrn.cum_loss})


class ForwardingWorker(WorkerNode):
    """`SingleLearner`: workers forward raw records to one learner at hub 0 (forced for HT and
    K-means, FlinkSpoke.scala:203-211).  The worker itself creates the hub (the reference relies
    on the out-of-tree worker for that: worker 0 never sends Create, FlinkSpoke.scala:220)."""
    protocol = "SingleLearner"

    def bind(self, node_id, network):
        super().bind(node_id, network)

    def receive_tuple(self, point):
        self.count += 1
        self.network.send(self.node_id, NodeId(NodeType.HUB, 0), RemoteCallIdentifier("forward"), point)

    def receive_query(self, response_id, payload):
        mean_buf, test = payload
        if response_id != -1 and self.node_id.node_id != 0:
            return   # one answer per query: the single learner lives at the hub
        self.network.send(self.node_id, NodeId(NodeType.HUB, 0), RemoteCallIdentifier("query"),
                          {"responseId": response_id, "test": test, "mean_buf": mean_buf})


class AsynchronousWorker(WorkerNode):
    protocol = "Asynchronous"

    def on_mini_batch(self):
        if not self.ready():
            return
        self.ensure_e()
        self.blocked = True
        self.push("push_delta", (self.flat - self.e) * self.scale(), self._extra())

    def _extra(self):
        return None

    def rpc_update_model(self, source, data):
        if self.install_shard(source, data):
            self.blocked = False


class SynchronousWorker(AsynchronousWorker):
    protocol = "Synchronous"


class SSPWorker(AsynchronousWorker):
    protocol = "SSP"

    def configure_worker(self, request):
        super().configure_worker(request)
        self.clock = 0
        self.staleness = cfg_int(self.tc, ["staleness", "s", "stalenessBound"], 3)
        return self

    def _extra(self):
        self.clock += 1
        return {"clock": self.clock}


class EASGDWorker(WorkerNode):
    protocol = "EASGD"

    def configure_worker(self, request):
        super().configure_worker(request)
        self.alpha = cfg_float(self.tc, ["alpha", "elasticity", "rho"], 0.1)
        self.tau = cfg_int(self.tc, ["tau", "commPeriod"], 1)
        self._mb = 0
        return self

    def on_mini_batch(self):
        if not self.ready():
            return
        self._mb += 1
        if self._mb % self.tau:
            return
        self.blocked = True
        self.push("elastic", self.flat.copy())

    def rpc_elastic_diff(self, source, data):
        self._pending[source.node_id] = data
        if len(self._pending) < self.H:
            return
        x = self.flat.copy()
        for h, (lo, hi) in enumerate(shard_bounds(len(x), self.H)):
            x[lo:hi] -= self._pending[h]["values"]
        self._pending = {}
        self.pipeline.learner.set_flat(x)
        self.blocked = False


class GMWorker(WorkerNode):
    """Geometric monitoring with a ball safe zone: ||w - e||^2 <= T."""
    protocol = "GM"

    def configure_worker(self, request):
        super().configure_worker(request)
        self.threshold = cfg_float(self.tc, ["threshold", "T", "gmThreshold"], 0.5)
        return self

    def on_mini_batch(self):
        if not self.ready():
            return
        self.ensure_e()
        d = self.flat - self.e
        if float(d @ d) > self.threshold:
            self.blocked = True
            self.network.send(self.node_id, NodeId(NodeType.HUB, 0), RemoteCallIdentifier("violation"), {})

    def rpc_send_model(self, source, data):
        if not self.ready():
            self.pipeline.learner.ensure_init(int(data["dim"])) if data.get("dim") else None
        if not self.ready():
            self.network.send(self.node_id, NodeId(NodeType.HUB, 0), RemoteCallIdentifier("no_model"), {})
            return
        self.ensure_e()
        self.blocked = True
        self.push("push_model", self.flat.copy())

    def rpc_update_model(self, source, data):
        if self.install_shard(source, data):
            self.blocked = False


class FGMWorker(WorkerNode):
    """Functional geometric monitoring; safe function phi(D) = ||D|| - sqrt(T), D = w - e."""
    protocol = "FGM"

    def configure_worker(self, request):
        super().configure_worker(request)
        self.threshold = cfg_float(self.tc, ["threshold", "T", "fgmThreshold"], 0.5)
        self.theta = math.sqrt(self.threshold) / 2.0
        self.zeta = -math.sqrt(self.threshold)
        self.sent = 0
        return self

    def phi(self) -> float:
        self.ensure_e()
        d = self.flat - self.e
        return float(math.sqrt(d @ d)) - math.sqrt(self.threshold)

    def on_mini_batch(self):
        if not self.ready():
            return
        ci = max(self.sent, int(math.floor((self.phi() - self.zeta) / self.theta))) if self.theta > 0 else self.sent
        if ci > self.sent:
            inc, self.sent = ci - self.sent, ci
            self.network.send(self.node_id, NodeId(NodeType.HUB, 0), RemoteCallIdentifier("increment"), {"inc": inc})

    def rpc_poll_phi(self, source, data):
        self.network.send(self.node_id, NodeId(NodeType.HUB, 0), RemoteCallIdentifier("phi"),
                          {"phi": self.phi() if self.ready() else -math.sqrt(self.threshold)})

    def rpc_new_subround(self, source, data):
        self.theta = float(data["theta"])
        self.zeta = self.phi() if self.ready() else -math.sqrt(self.threshold)
        self.sent = 0

    def rpc_send_model(self, source, data):
        if not self.ready():
            self.network.send(self.node_id, NodeId(NodeType.HUB, 0), RemoteCallIdentifier("no_model"), {})
            return
        self.ensure_e()
        self.blocked = True
        self.push("push_model", self.flat.copy())

    def rpc_update_model(self, source, data):
        if self.install_shard(source, data):
            self.blocked = False
            self.theta = math.sqrt(self.threshold) / 2.0
            self.zeta = -math.sqrt(self.threshold)
            self.sent = 0


# ============================================================================== parameter servers
class ParameterServerNode:
    """Base of the hub-side nodes: `mlParameterServers.MLParameterServer[_,_]` in the reference's out-of-tree API
    (accessors read at FlinkHub.scala:98-127: protocol statistics, fitted count, node id, learning curve, idx)."""
    protocol = "Asynchronous"

    def __init__(self):
        self.node_id: Optional[NodeId] = None
        self.network = None
        self.request: Optional[Request] = None
        self.global_shard: Optional[np.ndarray] = None
        self.stats = ProtocolStatistics(self.protocol)
        self.fitted = 0
        self.learning_curve: List[Tuple[float, int]] = []
        self.idx = 0
        self._asm: Dict[int, np.ndarray] = {}
        self._meta: Dict[int, dict] = {}
        self._base: Optional[np.ndarray] = None

    def configure_parameter_server(self, request: Request) -> "ParameterServerNode":
        self.request = request
        self.tc = request.trainingConfiguration or {}
        self.mini_batch = cfg_int(self.tc, ["miniBatchSize", "mini_batch_size", "batchSize"], 64)
        self.max_msg_params = cfg_int(self.tc, ["maxMsgParams"], 2000)
        self.stats = ProtocolStatistics(self.protocol)
        return self

    def bind(self, node_id, network):
        self.node_id, self.network = node_id, network

    # accessors the reference reads for statistics (FlinkHub.scala:98-127)
    def get_protocol_statistics(self) -> ProtocolStatistics:
        return self.stats

    def get_number_of_fitted_data(self) -> int:
        return self.fitted

    def get_learning_curve(self):
        return self.learning_curve

    def get_node_id(self):
        return self.node_id

    def get_idx(self) -> int:
        return self.idx

    def set_idx(self, idx: int) -> None:
        self.idx = int(idx)

    def n_workers(self) -> int:
        return self.network.describe().n_spokes

    def receive_msg(self, source: NodeId, rpc: RemoteCallIdentifier, data: Any) -> None:
        getattr(self, "rpc_" + rpc.op)(source, data)

    def assemble(self, source: NodeId, data: dict) -> Optional[np.ndarray]:
        """Reassemble the blocks of one worker's shard; returns it when the last block arrived."""
        w = source.node_id
        if w not in self._asm or len(self._asm[w]) != data["shard_len"]:
            self._asm[w] = np.zeros(data["shard_len"])
        v = data["values"]
        self._asm[w][data["offset"]:data["offset"] + len(v)] = v
        if "base" in data and self.global_shard is None:
            if self._base is None or len(self._base) != data["shard_len"]:
                self._base = np.zeros(data["shard_len"])
            self._base[data["offset"]:data["offset"] + len(v)] = data["base"]
        self.stats.num_of_blocks += 1
        self.stats.bytes_shipped += 8 * len(v)
        if not data["last"]:
            return None
        self._meta[w] = data
        return self._asm.pop(w)

    def reply(self, worker: NodeId, op: str, values: np.ndarray, total: int) -> None:
        self.stats.ship(len(values), max(1, math.ceil(len(values) / self.max_msg_params)))
        self.network.send(self.node_id, worker, RemoteCallIdentifier(op), {"values": values.copy(), "total": total})

    def broadcast(self, op: str, values: Optional[np.ndarray], total: int, extra: Optional[dict] = None) -> None:
        k = self.n_workers()
        dests = {NodeId(NodeType.SPOKE, i): RemoteCallIdentifier(op) for i in range(k)}
        payload = {"total": total}
        if values is not None:
            payload["values"] = values.copy()
            self.stats.ship(len(values) * k, k * max(1, math.ceil(len(values) / self.max_msg_params)), models=k)
        if extra:
            payload.update(extra)
        self.network.broadcast(self.node_id, dests, payload)

    def note_progress(self, n: int) -> None:
        self.fitted += n

    def rpc_report(self, source, data):
        self.fitted += int(data.get("fitted", 0))
        self.learning_curve.append((float(data.get("loss", 0.0)), self.fitted))

    def rpc_no_model(self, source, data):
        pass


class SimplePS(ParameterServerNode):
    protocol = "CentralizedTraining"


class AsynchronousParameterServer(ParameterServerNode):
    protocol = "Asynchronous"

    def _fold(self, shard: np.ndarray) -> None:
        if self.global_shard is None or len(self.global_shard) != len(shard):
            self.global_shard = self._base.copy() if self._base is not None and len(self._base) == len(shard) \
                else np.zeros_like(shard)
        self.global_shard += shard

    def rpc_push_delta(self, source, data):
        shard = self.assemble(source, data)
        if shard is None:
            return
        self._fold(shard)
        self.note_progress(self.mini_batch)
        self.reply(source, "update_model", self.global_shard, data["total"])


class SynchronousParameterServer(AsynchronousParameterServer):
    protocol = "Synchronous"

    def configure_parameter_server(self, request):
        super().configure_parameter_server(request)
        self._arrived: Dict[int, np.ndarray] = {}
        return self

    def rpc_push_delta(self, source, data):
        shard = self.assemble(source, data)
        if shard is None:
            return
        self._arrived[source.node_id] = shard
        self.note_progress(self.mini_batch)
        if len(self._arrived) < self.n_workers():
            return
        for s in self._arrived.values():   # deltas are pre-scaled by the workers (1/k or 1)
            self._fold(s)
        self._arrived = {}
        self.broadcast("update_model", self.global_shard, data["total"])


class SSPParameterServer(AsynchronousParameterServer):
    protocol = "SSP"

    def configure_parameter_server(self, request):
        super().configure_parameter_server(request)
        self.staleness = cfg_int(self.tc, ["staleness", "s", "stalenessBound"], 3)
        self.clocks: Dict[int, int] = {}
        self.deferred: List[Tuple[NodeId, int, int]] = []
        return self

    def min_clock(self) -> int:
        k = self.n_workers()
        return min((self.clocks.get(i, 0) for i in range(k)), default=0)

    def rpc_push_delta(self, source, data):
        shard = self.assemble(source, data)
        if shard is None:
            return
        self._fold(shard)
        self.note_progress(self.mini_batch)
        c = int(data.get("clock", 0))
        self.clocks[source.node_id] = c
        self.deferred.append((source, c, data["total"]))
        still = []
        for (w, cw, total) in self.deferred:
            if cw - self.min_clock() <= self.staleness:
                self.reply(w, "update_model", self.global_shard, total)
            else:
                still.append((w, cw, total))
        self.deferred = still


class EASGDParameterServer(ParameterServerNode):
    protocol = "EASGD"

    def configure_parameter_server(self, request):
        super().configure_parameter_server(request)
        self.alpha = cfg_float(self.tc, ["alpha", "elasticity", "rho"], 0.1)
        return self

    def rpc_elastic(self, source, data):
        shard = self.assemble(source, data)
        if shard is None:
            return
        if self.global_shard is None or len(self.global_shard) != len(shard):
            self.global_shard = shard.copy()   # the centre variable starts at the first model seen
        diff = self.alpha * (shard - self.global_shard)
        self.global_shard += diff
        self.note_progress(self.mini_batch)
        self.stats.ship(len(diff))
        self.network.send(self.node_id, source, RemoteCallIdentifier("elastic_diff"),
                          {"values": diff, "total": data["total"]})


class GMParameterServer(ParameterServerNode):
    """Hub 0 coordinates (violation -> collect -> average -> broadcast); every hub averages its shard."""
    protocol = "GM"

    def configure_parameter_server(self, request):
        super().configure_parameter_server(request)
        self._arrived: Dict[int, np.ndarray] = {}
        self._missing = 0
        self.collecting = False
        self.rounds = 0
        return self

    def rpc_violation(self, source, data):
        if self.collecting:
            return
        self.collecting = True
        self.broadcast("send_model", None, 0)

    def rpc_no_model(self, source, data):
        self._missing += 1
        self._maybe_finish(None)

    def rpc_push_model(self, source, data):
        shard = self.assemble(source, data)
        if shard is None:
            return
        self._arrived[source.node_id] = shard
        self._total = data["total"]
        self._maybe_finish(data)

    def _maybe_finish(self, data):
        if len(self._arrived) + self._missing < self.n_workers() or not self._arrived:
            return
        self.global_shard = np.mean(list(self._arrived.values()), axis=0)
        self.note_progress(0)
        self._arrived, self._missing, self.collecting = {}, 0, False
        self.rounds += 1
        self.broadcast("update_model", self.global_shard, self._total)


class FGMParameterServer(GMParameterServer):
    protocol = "FGM"

    def configure_parameter_server(self, request):
        super().configure_parameter_server(request)
        self.threshold = cfg_float(self.tc, ["threshold", "T", "fgmThreshold"], 0.5)
        self.eps_psi = cfg_float(self.tc, ["epsPsi", "eps_psi"], 0.01)
        self.counter = 0
        self.polling = False
        self._phis: Dict[int, float] = {}
        self.subrounds = 0
        return self

    def rpc_increment(self, source, data):
        if self.collecting or self.polling:
            return
        self.counter += int(data["inc"])
        if self.counter > self.n_workers():
            self.polling = True
            self._phis = {}
            self.broadcast("poll_phi", None, 0)

    def rpc_phi(self, source, data):
        self._phis[source.node_id] = float(data["phi"])
        k = self.n_workers()
        if len(self._phis) < k:
            return
        psi = sum(self._phis.values())
        self.polling = False
        self.counter = 0
        phi0 = -math.sqrt(self.threshold)
        if psi >= self.eps_psi * k * phi0:
            self.collecting = True           # round ends: collect models, average, broadcast
            self.broadcast("send_model", None, 0)
        else:
            self.subrounds += 1
            self.broadcast("new_subround", None, 0, {"theta": -psi / (2.0 * k)})


class CentralizedMLServer(ParameterServerNode):
    """`SingleLearner` hub: trains the only pipeline on forwarded records, answers queries and
    forecasts from the hub side (FlinkNetwork.scala:255, FlinkHub.scala:128-153)."""
    protocol = "SingleLearner"

    def configure_parameter_server(self, request):
        super().configure_parameter_server(request)
        self.pipeline = OraclePipeline(request)
        return self

    def rpc_forward(self, source, point):
        from ..stream.points import ForecastingPoint
        if isinstance(point, ForecastingPoint):
            pred = self.pipeline.predict(point.numerical)
            self.network.send(self.node_id, None, RemoteCallIdentifier("prediction"),
                              Prediction(self.request.id, point.data_instance, pred))
            return
        loss = self.pipeline.fit(point.numerical, point.target)
        self.fitted += 1
        if self.fitted % self.mini_batch == 0:
            self.learning_curve.append((self.pipeline.window_loss / max(1, self.pipeline.window_n), self.fitted))

    def rpc_query(self, source, data):
        X, y = data["test"]
        resp = self.pipeline.query_response(data["responseId"], self.protocol, (X, y))
        if data["responseId"] == -1:
            resp.cumulativeLoss = float(data.get("mean_buf", 0.0))
        self.network.send(self.node_id, None, RemoteCallIdentifier("query_response"), resp)


WORKERS = {"CentralizedTraining": SingleWorker, "SingleLearner": ForwardingWorker,
           "Asynchronous": AsynchronousWorker, "Synchronous": SynchronousWorker, "SSP": SSPWorker,
           "EASGD": EASGDWorker, "GM": GMWorker, "FGM": FGMWorker}
SERVERS = {"CentralizedTraining": SimplePS, "SingleLearner": CentralizedMLServer,
           "Asynchronous": AsynchronousParameterServer, "Synchronous": SynchronousParameterServer,
           "SSP": SSPParameterServer, "EASGD": EASGDParameterServer, "GM": GMParameterServer,
           "FGM": FGMParameterServer}


MLParameterServer = ParameterServerNode      # the reference's name for the base class (FlinkHub.scala:98)
NodeInstance = (WorkerNode, ParameterServerNode)   # `NodeInstance[_,_]`, base of every worker / PS (NodeGenerator.scala:8-9)


class MLNodeGenerator:
    """protocol string -> (worker class, PS class); unknown / missing / failing -> Asynchronous
    (utils/generators/MLNodeGenerator.scala:20-76)."""

    def __init__(self, max_msg_params: int = 2000):
        self.max_msg_params = max_msg_params

    @staticmethod
    def protocol_of(request: Request) -> str:
        tc = request.trainingConfiguration or {}
        p = tc.get("protocol")
        return p if p in WORKERS else "Asynchronous"

    def generate_spoke_node(self, request: Request) -> WorkerNode:
        try:
            return WORKERS[self.protocol_of(request)](self.max_msg_params).configure_worker(request)
        except Exception:
            return AsynchronousWorker(self.max_msg_params).configure_worker(request)

    def generate_hub_node(self, request: Request) -> ParameterServerNode:
        try:
            return SERVERS[self.protocol_of(request)]().configure_parameter_server(request)
        except Exception:
            return AsynchronousParameterServer().configure_parameter_server(request)
