"""Request validation and pipeline registry (the single-instance `PipelineMap` operator).

Reference: utils/parsers/requestStream/PipelineMap.scala - whitelists :66-69, Create :31-34,
Update :35-36, Query :37-42 (HT / K-means queries go to worker 0 only), Delete :43-46,
broadcast :54-57.  Live pipeline ids are keyed state there (:59-64); here a dict that is part of
the checkpoint.
"""
from __future__ import annotations

from typing import Dict, List

from ..messages import ControlMessage, NodeId
from .pojos import LEARNERS, PREPROCESSORS, Request

SINGLE_LEARNER_ONLY = ("HT", "K-means")


class PipelineMap:
    def __init__(self, parallelism: int, distributed_kmeans: bool = False):
        self.parallelism = parallelism
        self.nodes: Dict[int, Request] = {}
        # BASELINE.json config 4 asks for k-means under SSP on 8 workers, which the reference
        # cannot do (FlinkSpoke.scala:207-209); opting in lifts the worker-0-only query rule.
        self.distributed_kmeans = distributed_kmeans

    def _valid_names(self, req: Request) -> bool:
        if req.learner is None or req.learner.name not in LEARNERS:
            return False
        for p in req.preProcessors or []:
            if p.name not in PREPROCESSORS:
                return False
        return True

    def _broadcast(self, req: Request) -> List[ControlMessage]:
        return [ControlMessage(req.id, None, None, NodeId(None, i), None, req) for i in range(self.parallelism)]

    def _send(self, req: Request, worker: int = 0) -> List[ControlMessage]:
        return [ControlMessage(req.id, None, None, NodeId(None, worker), None, req)]

    def flat_map(self, req: Request) -> List[ControlMessage]:
        if not req.is_valid():
            return []
        kind = req.request
        if kind == "Create":
            if req.id in self.nodes or not self._valid_names(req):
                return []
            self.nodes[req.id] = req
            return self._broadcast(req)
        if kind == "Update":
            if req.id not in self.nodes:
                return []
            return self._broadcast(req)
        if kind == "Query":
            if req.id not in self.nodes:
                return []
            name = self.nodes[req.id].learner.name
            single = name == "HT" or (name == "K-means" and not self.distributed_kmeans)
            return self._send(req, 0) if single else self._broadcast(req)
        if kind == "Delete":
            if req.id not in self.nodes:
                return []
            del self.nodes[req.id]
            return self._broadcast(req)
        return []

    # checkpoint support (PipelineMap.scala:59-64)
    def state_dict(self) -> dict:
        return {"nodes": {k: v.to_dict() for k, v in self.nodes.items()}, "parallelism": self.parallelism}

    def load_state_dict(self, sd: dict) -> None:
        self.nodes = {int(k): Request.from_dict(v) for k, v in sd["nodes"].items()}
