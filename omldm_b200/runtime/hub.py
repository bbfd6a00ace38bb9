"""Parameter-server ("hub") operator of the host runtime, keyed by "networkId_hubId".

Reference: operators/hub/FlinkHub.scala - `processElement` :54-162 (Create :62, Delete :65,
cache of early worker messages :70-71 with cap 20 000, delivery :73-87, per-message Statistics in
test mode :88-157), `generateHub` :166-195; state/StateAccumulators.scala:128-173
(`DataAggregateFunction` FIFO cache, `NodeAggregateFunction.add` = deliver to the wrapped node).
"""
from __future__ import annotations

from collections import deque
from typing import Any, Callable, Deque, Optional

from ..api.pojos import Statistics
from ..messages import HubMessage, NodeId, NodeType, SpokeMessage
from ..parallel.host_protocols import MLNodeGenerator, cfg
from ..parallel.network import GenericWrapper, IntWrapper, OperatorNetwork
from ..utils.config import HUB_CACHE_CAP


class Hub:
    def __init__(self, network_id: int, hub_id: int, parallelism: int, *, max_msg_params: int, test: bool,
                 collect: Callable[[HubMessage], None], side_output: Callable[[Any], None],
                 stats_output: Callable[[int, int, Statistics], None]):
        self.network_id, self.hub_id = network_id, hub_id
        self.parallelism = parallelism
        self.test = test
        self.collect, self.side_output, self.stats_output = collect, side_output, stats_output
        self.node_factory = MLNodeGenerator(max_msg_params)
        self.state: Optional[GenericWrapper] = None
        self.cache: Deque[SpokeMessage] = deque()

    def process_element(self, msg: SpokeMessage) -> None:
        if msg.request is not None:
            kind = msg.request.request
            if kind == "Create":
                self.generate_hub(msg)
            elif kind == "Update":
                pass  # no-op (FlinkHub.scala:63)
            elif kind == "Delete":
                self.state = None
                self.cache.clear()
            return
        if self.state is None:
            if len(self.cache) >= HUB_CACHE_CAP:
                self.cache.popleft()
            self.cache.append(msg)
            return
        while self.cache:
            self._deliver(self.cache.popleft())
        self._deliver(msg)

    def _deliver(self, msg: SpokeMessage) -> None:
        self.state.receive_msg(msg.source, msg.operation, msg.data)
        if self.test:
            self.stats_output(self.network_id, self.hub_id, self.statistics())

    def statistics(self) -> Statistics:
        node = self.state.get_node()
        ps = node.get_protocol_statistics()
        curve = node.get_learning_curve()
        return Statistics(pipeline=self.network_id, protocol=ps.protocol, models=ps.models_shipped,
                          bytes=ps.bytes_shipped, blocks=ps.num_of_blocks,
                          fitted=node.get_number_of_fitted_data(),
                          learningCurve=[c[0] for c in curve], lcx=[c[1] for c in curve])

    def generate_hub(self, msg: SpokeMessage) -> None:
        if self.state is not None:
            return
        req = msg.request
        hub_par = int(cfg(req.trainingConfiguration, ["HubParallelism", "hubParallelism"], 1))
        if self.parallelism == 1:   # FlinkHub.scala:185-190
            if req.trainingConfiguration is None:
                req.trainingConfiguration = {}
            if req.trainingConfiguration.get("protocol") != "SingleLearner":
                req.trainingConfiguration["protocol"] = "CentralizedTraining"
        network = OperatorNetwork(NodeType.HUB, self.network_id, IntWrapper(self.parallelism), hub_par,
                                  self.collect, self.side_output)
        node = self.node_factory.generate_hub_node(req)
        self.state = GenericWrapper(NodeId(NodeType.HUB, self.hub_id), node, network)

    def snapshot_state(self) -> dict:
        node = self.state.get_node() if self.state else None
        return {"network_id": self.network_id, "hub_id": self.hub_id,
                "request": node.request.to_dict() if node else None,
                "global_shard": None if node is None or node.global_shard is None else node.global_shard.copy(),
                "fitted": node.fitted if node else 0,
                "cache": list(self.cache)}
