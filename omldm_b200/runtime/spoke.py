"""Worker ("spoke") operator of the host runtime.

Reference: operators/spoke/FlinkSpoke.scala - `processElement1` :63-90, `handleData` :92-107
(2-in-10 holdout, evict-then-train), `processElement2` :117-175, `createWrapper` :177-224
(protocol rewrites :202-216, worker 0 creates the hubs :220-222), snapshot/restore :233-334,
`checkParallelism` :345-348; base class operators/spoke/SpokeLogic.scala:28-50 (state map,
100 000-record buffer, 10 000-request buffer).

Fixes relative to the reference: restored pipelines are actually reinstated, and
under SingleLearner the hub is created by worker 0 like for every other protocol.
"""
from __future__ import annotations

import copy
from typing import Any, Callable, Dict, List

import numpy as np

from ..api.pojos import Request
from ..messages import ControlMessage, NodeId, NodeType, SpokeMessage
from ..oracle.pipeline import HoldoutSampler
from ..parallel.host_protocols import MLNodeGenerator, cfg
from ..parallel.network import BufferingWrapper, DataSet, IntWrapper, OperatorNetwork
from ..stream.points import ForecastingPoint, TrainingPoint, UsablePoint
from ..utils.config import RECORD_BUFFER_CAP, REQUEST_BUFFER_CAP, TICK_EVERY


class Spoke:
    def __init__(self, worker_id: int, parallelism: int, *, max_msg_params: int, test: bool,
                 test_set_size: int, collect: Callable[[SpokeMessage], None],
                 side_output: Callable[[Any], None], distributed_kmeans: bool = False):
        self.worker_id = worker_id
        self.expected_parallelism = parallelism
        self.spoke_parallelism = IntWrapper(parallelism)
        self.test = test
        self.collect, self.side_output = collect, side_output
        self.node_factory = MLNodeGenerator(max_msg_params)
        self.state: Dict[int, BufferingWrapper] = {}
        self.record_buffer: DataSet[UsablePoint] = DataSet(RECORD_BUFFER_CAP)
        self.request_buffer: DataSet[ControlMessage] = DataSet(REQUEST_BUFFER_CAP)
        self.holdout = HoldoutSampler(test_set_size)
        self.count = 0
        self.distributed_kmeans = distributed_kmeans

    # ------------------------------------------------------------------ data (processElement1)
    def process_element1(self, data: UsablePoint) -> None:
        while self.request_buffer.non_empty():
            self._create(self.request_buffer.pop())
        if not self.state:
            self.record_buffer.append(data)
        else:
            while self.record_buffer.non_empty():
                self.handle_data(self.record_buffer.pop())
            self.handle_data(data)
        if self.test:
            self.count += 1
            if self.count % TICK_EVERY == 0:  # "alive" tick (FlinkSpoke.scala:83-89)
                self.collect(SpokeMessage(-1, None, None, None, None, None))

    def handle_data(self, data: UsablePoint) -> None:
        if isinstance(data, TrainingPoint):
            routed = self.holdout.route(data.numerical, data.target)
            if routed is None:
                return
            pt = data if routed[0] is data.numerical else TrainingPoint(routed[0], target=routed[1])
            for w in self.state.values():
                w.receive_tuple([pt])
        elif isinstance(data, ForecastingPoint):
            for w in self.state.values():
                w.receive_tuple([data])

    # ------------------------------------------------------------------ control (processElement2)
    def process_element2(self, msg: ControlMessage) -> None:
        net = msg.network_id
        if msg.request is None:
            if net == -1:  # termination probe (FlinkSpoke.scala:136-138)
                for w in self.state.values():
                    w.receive_query(-1, (w.get_mean_buffer_size() / self.spoke_parallelism.get_int(),
                                         self.holdout.arrays()))
                return
            if net in self.state:  # hub RPC delivery (:127-130)
                self.state[net].receive_msg(msg.source, msg.operation, msg.data)
                for other, w in self.state.items():
                    if other != net:
                        w.toggle()
            return
        kind = msg.request.request
        if kind == "Create":
            self._create(msg)
        elif kind == "Update":
            pass  # accepted and routed, but a no-op (FlinkSpoke.scala:158)
        elif kind == "Query":
            if net in self.state:
                self.state[net].receive_query(msg.request.requestId, (0, self.holdout.arrays()))
        elif kind == "Delete":
            self.state.pop(net, None)

    def check_parallelism(self, observed: int) -> None:
        """Elasticity hook: raise the shared worker count if the job was rescaled up."""
        if observed > self.spoke_parallelism.get_int():
            self.spoke_parallelism.set_int(observed)

    def _create(self, msg: ControlMessage) -> None:
        req = copy.deepcopy(msg.request)
        net = req.id
        if net in self.state:
            return
        tc = req.trainingConfiguration if req.trainingConfiguration is not None else {}
        req.trainingConfiguration = tc
        hub_par = int(cfg(tc, ["HubParallelism", "hubParallelism"], 1))
        name = req.learner.name
        # protocol rewrites (FlinkSpoke.scala:202-216)
        if name == "HT" or (name == "K-means" and not self.distributed_kmeans):
            tc["protocol"] = "SingleLearner"
        elif self.spoke_parallelism.get_int() == 1:
            tc["protocol"] = "CentralizedTraining"
        if tc.get("protocol") in ("SingleLearner", "CentralizedTraining"):
            hub_par = 1
        tc["HubParallelism"] = hub_par
        network = OperatorNetwork(NodeType.SPOKE, net, self.spoke_parallelism, hub_par, self.collect,
                                  self.side_output)
        node = self.node_factory.generate_spoke_node(req)
        self.state[net] = BufferingWrapper(NodeId(NodeType.SPOKE, self.worker_id), node, network)
        if self.worker_id == 0:  # worker 0 asks every hub shard to come up (:220-222)
            for h in range(hub_par):
                self.collect(SpokeMessage(net, None, None, NodeId(NodeType.HUB, h), None, req))

    # ------------------------------------------------------------------ checkpoint
    def snapshot_state(self) -> dict:
        return {
            "worker_id": self.worker_id,
            "pipelines": {net: {"request": w.node.request.to_dict(),
                                "flat": None if not w.node.ready() else w.node.flat.copy(),
                                "dim": w.node.pipeline.learner.d,
                                "fitted": w.node.pipeline.learner.fitted,
                                "cum_loss": w.node.pipeline.learner.cum_loss,
                                "count": w.node.count}
                          for net, w in self.state.items()},
            "test_set": list(self.holdout.test_set), "holdout_count": self.holdout.count,
            "record_buffer": list(self.record_buffer.data_buffer),
            "request_buffer": list(self.request_buffer.data_buffer),
            "count": self.count,
        }

    def initialize_state(self, snaps: List[dict]) -> None:
        """Restore from one or several old subtask snapshots (rescaling merges them)."""
        for i, s in enumerate(snaps):
            for net, p in s["pipelines"].items():
                if net not in self.state:
                    self._create(ControlMessage(net, None, None, None, None, Request.from_dict(p["request"])))
                    node = self.state[net].node
                    if p["flat"] is not None:
                        node.pipeline.learner.ensure_init(p["dim"])
                        node.pipeline.learner.set_flat(p["flat"])
                        node.e = np.array(p["flat"], dtype=np.float64)
                    node.pipeline.learner.fitted = p["fitted"]
                    node.pipeline.learner.cum_loss = p["cum_loss"]
                    node.count = p["count"]
            for pt in s["test_set"]:
                self.holdout.test_set.append(pt)
            self.holdout.count += s["holdout_count"] if i == 0 else 0
            other = DataSet(RECORD_BUFFER_CAP)
            other.data_buffer.extend(s["record_buffer"])
            self.record_buffer.merge([other])
            for r in s["request_buffer"]:
                self.request_buffer.append(r)
            self.count = max(self.count, s["count"])
        while len(self.holdout.test_set) > self.holdout.capacity:
            self.holdout.test_set.popleft()
