"""The host (CPU, fp64) job: the whole dataflow of the reference in one process.

Reference: Job.scala:35-108,110-168 (sources, PipelineMap at parallelism 1, sinks) and
job/FlinkLearning.scala:33-150 (hub->worker fan-out :65-75, identity partitioning :78-80, spoke
:86-88, keyBy(net_hub) + hub :91-95, predictions :98-101, query responses :104-109, performance
:112-144, feedback :147).  Kafka topics become in-process queues; the psMessages feedback topic
becomes the in-flight message queue.  This engine is the correctness oracle and the timed CPU
baseline that stands in for the reference's "local CPU mini-cluster" (BASELINE.json config 1);
the GPU engine (`omldm_b200/engine.py`) offers the same request/response surface.
"""
from __future__ import annotations

import time
from collections import deque
from typing import Any, Callable, Deque, Dict, Iterable, List, Optional, Union

from ..api.pojos import DataInstance, JobStatistics, Prediction, QueryResponse, Request, Statistics
from ..api.registry import PipelineMap
from ..api.response_merger import ResponseConstructor
from ..messages import ControlMessage, HubMessage, NodeId, NodeType, SpokeMessage
from ..stream.parsers import parse_data_instance, parse_request
from ..stream.points import UsablePoint, data_point_from_instance
from ..utils.config import JobParameters
from ..utils.partitioners import hub_key, identity_partitioner
from .hub import Hub
from .spoke import Spoke
from .statistics import JobTerminated, StatisticsOperator, job_terminator, performance_writer


class HostJob:
    def __init__(self, params: Optional[JobParameters] = None, clock: Optional[Callable[[], float]] = None):
        self.params = params or JobParameters()
        P = self.params.parallelism
        self.clock = clock or (lambda: time.monotonic() * 1000.0)
        self.pipeline_map = PipelineMap(P, self.params.distributedKMeans)
        self.response_constructor = ResponseConstructor(P)
        self.stats_op = StatisticsOperator(self.params.jobName, P, self.params.timeout, self.params.testSetSize)
        self.in_flight: Deque[Any] = deque()
        self.responses: List[str] = []
        self.predictions: List[str] = []
        self.performance: List[str] = []
        self.job_statistics: Optional[JobStatistics] = None
        self.spokes = [Spoke(i, P, max_msg_params=self.params.maxMsgParams, test=self.params.test,
                             test_set_size=self.params.testSetSize, collect=self.in_flight.append,
                             side_output=self._side_output,
                             distributed_kmeans=self.params.distributedKMeans) for i in range(P)]
        self.hubs: Dict[str, Hub] = {}
        self._rr = 0
        self.terminated = False
        self.records_in = 0

    # ------------------------------------------------------------------ inputs
    def submit_request(self, req: Union[str, bytes, dict, Request]) -> None:
        r = parse_request(req)
        if r is None:
            return
        if r.request == "Query" and r.id in self.pipeline_map.nodes:
            created = self.pipeline_map.nodes[r.id]
            name = created.learner.name
            single = name == "HT" or (name == "K-means" and not self.params.distributedKMeans) or \
                (created.trainingConfiguration or {}).get("protocol") == "SingleLearner"
            if single:
                self.response_constructor.expect(r.requestId, 1)
        for cm in self.pipeline_map.flat_map(r):
            self.spokes[identity_partitioner(cm.destination.node_id, len(self.spokes))].process_element2(cm)
        if r.request == "Delete":
            for k in [k for k in self.hubs if k.split("_")[0] == str(r.id)]:
                del self.hubs[k]
        self._drain()

    def submit_data(self, rec: Union[str, bytes, dict, DataInstance, UsablePoint], worker: Optional[int] = None) -> None:
        pt = rec if isinstance(rec, UsablePoint) else None
        if pt is None:
            di = parse_data_instance(rec)
            if di is None:
                return
            pt = data_point_from_instance(di)
            if pt is None:
                return
        w = self._rr % len(self.spokes) if worker is None else worker
        self._rr += 1
        self.records_in += 1
        self.spokes[w].process_element1(pt)
        self._drain()

    # ------------------------------------------------------------------ message plumbing
    def _drain(self) -> None:
        while self.in_flight:
            m = self.in_flight.popleft()
            if isinstance(m, SpokeMessage):
                self._to_hub(m)
            elif isinstance(m, HubMessage):
                self._feedback(m)

    def _to_hub(self, m: SpokeMessage) -> None:
        now = self.clock()
        if m.network_id == -1:   # alive tick -> statistics operator only (FlinkLearning.scala:112-135)
            self.stats_op.on_activity(now)
            return
        key = hub_key(m.network_id, m.destination.node_id)
        hub = self.hubs.get(key)
        if hub is None:
            hub = Hub(m.network_id, m.destination.node_id, self.params.parallelism,
                      max_msg_params=self.params.maxMsgParams, test=self.params.test,
                      collect=self.in_flight.append, side_output=self._side_output,
                      stats_output=self._hub_stats)
            self.hubs[key] = hub
        hub.process_element(m)

    def _hub_stats(self, net: int, hub: int, st: Statistics) -> None:
        self.stats_op.on_activity(self.clock(), net, hub, st)

    def _feedback(self, m: HubMessage) -> None:
        if m.network_id == -1:   # termination probe to every worker (FlinkLearning.scala:68-70)
            for i, s in enumerate(self.spokes):
                s.process_element2(ControlMessage(-1, None, None, NodeId(NodeType.SPOKE, i), None, None))
            return
        for cm in m.fan_out():
            self.spokes[identity_partitioner(cm.destination.node_id, len(self.spokes))].process_element2(cm)

    def _side_output(self, obj: Any) -> None:
        if isinstance(obj, Prediction):
            self.predictions.append(str(obj))
            return
        if isinstance(obj, QueryResponse):
            if obj.responseId == -1:   # FlinkLearning.scala:115-132
                st = Statistics(pipeline=obj.mlpId, meanBufferSize=float(obj.cumulativeLoss or 0.0),
                                fitted=int(obj.dataFitted or 0), score=float(obj.score or 0.0))
                js = self.stats_op.on_terminate(st, len(self.pipeline_map.nodes))
                if js is not None:
                    self._finish(js)
                return
            for merged in self.response_constructor.flat_map(obj):
                if merged.responseId >= 0:
                    self.responses.append(str(merged))

    def _finish(self, js: JobStatistics) -> None:
        if js.durationMs > 0:
            tot = sum(s.fitted for s in js.statistics)
            js.extra["recordsPerSec"] = tot / (js.durationMs / 1000.0)
            js.extra["syncBytesPerSec"] = sum(s.bytes for s in js.statistics) / (js.durationMs / 1000.0)
        self.job_statistics = js
        self.performance.append(performance_writer(js))
        self.terminated = True

    # ------------------------------------------------------------------ time / termination
    def advance_time(self, now_ms: Optional[float] = None) -> None:
        """Fire the quiescence timer if the stream has been quiet for `timeout` ms."""
        now = self.clock() if now_ms is None else now_ms
        if self.stats_op.on_time(now):
            self.in_flight.append(HubMessage())   # network id -1
            self._drain()

    def end_of_stream(self) -> Optional[JobStatistics]:
        """Convenience for batch runs: behave as if the quiescence timeout had elapsed."""
        if not self.params.test or self.terminated:
            return self.job_statistics
        if self.stats_op.timer is None:
            self.stats_op.timer = self.clock()
        self.advance_time(self.stats_op.timer + 1)
        return self.job_statistics

    def run(self, requests: Iterable[Any] = (), data: Iterable[Any] = (), terminate: bool = True) -> Optional[JobStatistics]:
        for r in requests:
            self.submit_request(r)
        for d in data:
            self.submit_data(d)
        js = self.end_of_stream() if terminate else None
        if js is not None and terminate:
            try:
                job_terminator(self.performance[-1])   # the reference's self-shutdown path
            except JobTerminated:
                pass
        return js

    # ------------------------------------------------------------------ checkpoint (utils/Checkpointing.scala)
    def snapshot(self) -> dict:
        return {"pipeline_map": self.pipeline_map.state_dict(),
                "spokes": [s.snapshot_state() for s in self.spokes],
                "hubs": {k: h.snapshot_state() for k, h in self.hubs.items()},
                "rr": self._rr, "records_in": self.records_in}

    def restore(self, snap: dict) -> None:
        self.pipeline_map.load_state_dict(snap["pipeline_map"])
        old = snap["spokes"]
        P = len(self.spokes)
        for i, s in enumerate(self.spokes):   # rescaling: old subtask j goes to new subtask j % P
            s.initialize_state([o for j, o in enumerate(old) if j % P == i])
        self._drain()
        for k, hs in snap["hubs"].items():
            h = self.hubs.get(k)
            if h is not None and h.state is not None and hs["global_shard"] is not None:
                h.state.get_node().global_shard = hs["global_shard"].copy()
                h.state.get_node().fitted = hs["fitted"]
        self._rr, self.records_in = snap["rr"], snap["records_in"]
