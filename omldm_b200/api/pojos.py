"""JSON control-API objects (the reference's out-of-tree `ControlAPI` POJOs).

Only call contracts are provable from the reference tree, so every class documents the call
site that fixes its members:

* Request            PipelineMap.scala:22-46, FlinkSpoke.scala:142-171, FlinkHub.scala:61-66
* LearnerPOJO        PipelineMap.scala:23,38, FlinkNetwork.scala:160-206
* PreprocessorPOJO   PipelineMap.scala:26-27
* DataInstance       DataPointParser.scala:17-46, DataInstanceParser.scala:15-17
* QueryResponse      FlinkNetwork.scala:196-230, ResponseConstructor.scala:34-53
* Prediction         FlinkNetwork.scala:250,255
* Statistics         FlinkHub.scala:118-127, StateAccumulators.scala:64-105
* JobStatistics      StatisticsOperator.scala:111-126

All objects serialise to the same field names the reference's Jackson mapping uses
(camelCase bean properties), so existing request producers / response consumers keep working.
"""
from __future__ import annotations

import json
import math
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

from .compat import CountableSerial

REQUEST_TYPES = ("Create", "Update", "Query", "Delete")
LEARNERS = ("PA", "RegressorPA", "ORR", "SVM", "MultiClassPA", "K-means", "NN", "HT")
PREPROCESSORS = ("PolynomialFeatures", "StandardScaler", "MinMaxScaler")
OPERATIONS = ("training", "forecasting")


def _clean(o: Any) -> Any:
    """Make numpy / torch scalars and arrays JSON-serialisable."""
    if o is None or isinstance(o, (bool, int, str)):
        return o
    if isinstance(o, float):
        return o if math.isfinite(o) else None
    if isinstance(o, dict):
        return {str(k): _clean(v) for k, v in o.items()}
    if isinstance(o, (list, tuple)):
        return [_clean(v) for v in o]
    if hasattr(o, "tolist"):
        return _clean(o.tolist())
    if hasattr(o, "to_dict"):
        return o.to_dict()
    return str(o)


def _sizeof(o: Any) -> int:
    """Byte accounting in the spirit of `CountableSerial.getSize` (FlinkMessage.scala:16-23)."""
    if o is None:
        return 0
    if isinstance(o, bool):
        return 1
    if isinstance(o, int):
        return 8
    if isinstance(o, float):
        return 8
    if isinstance(o, str):
        return len(o.encode())
    if isinstance(o, dict):
        return sum(_sizeof(k) + _sizeof(v) for k, v in o.items())
    if isinstance(o, (list, tuple)):
        return sum(_sizeof(v) for v in o)
    if hasattr(o, "nbytes"):
        return int(o.nbytes)
    if isinstance(o, CountableSerial):      # anything with get_size(): POJOs, messages, parameter containers
        return int(o.get_size())
    return 8


@dataclass
class LearnerPOJO:
    name: str = ""
    hyperParameters: Optional[Dict[str, Any]] = None
    parameters: Optional[Dict[str, Any]] = None
    dataStructure: Optional[Dict[str, Any]] = None

    @classmethod
    def from_dict(cls, d: Optional[dict]) -> Optional["LearnerPOJO"]:
        if d is None:
            return None
        return cls(d.get("name", ""), d.get("hyperParameters"), d.get("parameters"),
                   d.get("dataStructure"))

    def to_dict(self) -> dict:
        return _clean({"name": self.name, "hyperParameters": self.hyperParameters,
                       "parameters": self.parameters, "dataStructure": self.dataStructure})

    def get_size(self) -> int:
        return _sizeof(self.name) + _sizeof(self.hyperParameters) + _sizeof(self.parameters) + \
            _sizeof(self.dataStructure)


@dataclass
class PreprocessorPOJO(LearnerPOJO):
    pass


@dataclass
class Request:
    id: int = -1
    request: str = ""
    requestId: Optional[int] = None
    learner: Optional[LearnerPOJO] = None
    preProcessors: Optional[List[PreprocessorPOJO]] = None
    trainingConfiguration: Optional[Dict[str, Any]] = None
    metadata: Optional[Dict[str, Any]] = None

    # -- validation: `Request.isValid` (RequestParser.scala:14)
    def is_valid(self) -> bool:
        if not isinstance(self.id, int) or self.id < 0:
            return False
        if self.request not in REQUEST_TYPES:
            return False
        if self.request == "Create":
            if self.learner is None or not self.learner.name:
                return False
        if self.request == "Query" and self.requestId is None:
            return False
        return True

    @classmethod
    def from_dict(cls, d: dict) -> "Request":
        pps = d.get("preProcessors", d.get("preprocessors"))
        return cls(
            id=d.get("id", -1), request=d.get("request", ""), requestId=d.get("requestId"),
            learner=LearnerPOJO.from_dict(d.get("learner")),
            preProcessors=[PreprocessorPOJO.from_dict(p) for p in pps] if pps is not None else None,
            trainingConfiguration=d.get("trainingConfiguration", d.get("training_configuration")),
        )

    @classmethod
    def from_json(cls, s: str | bytes) -> "Request":
        return cls.from_dict(json.loads(s))

    def to_dict(self) -> dict:
        return _clean({
            "id": self.id, "request": self.request, "requestId": self.requestId,
            "learner": self.learner.to_dict() if self.learner else None,
            "preProcessors": [p.to_dict() for p in self.preProcessors] if self.preProcessors is not None else None,
            "trainingConfiguration": self.trainingConfiguration,
        })

    def to_json(self) -> str:
        return json.dumps(self.to_dict())

    def set_metadata(self, topic, partition, key, offset, timestamp) -> None:
        # RequestDeserializer.scala:25-30
        self.metadata = {"topic": topic, "partition": partition, "key": key, "offset": offset,
                         "timestamp": timestamp}

    def get_size(self) -> int:
        return 8 + _sizeof(self.request) + 8 + (self.learner.get_size() if self.learner else 0) + \
            sum(p.get_size() for p in (self.preProcessors or [])) + _sizeof(self.trainingConfiguration)


@dataclass
class DataInstance:
    numericalFeatures: Optional[List[float]] = None
    discreteFeatures: Optional[List[int]] = None
    categoricalFeatures: Optional[List[str]] = None
    target: Optional[float] = None
    operation: str = "training"
    id: Optional[int] = None
    metadata: Optional[Dict[str, Any]] = None

    def is_valid(self) -> bool:
        if self.operation not in OPERATIONS:
            return False
        has_any = any(f for f in (self.numericalFeatures, self.discreteFeatures, self.categoricalFeatures))
        return bool(has_any)

    @classmethod
    def from_dict(cls, d: dict) -> "DataInstance":
        return cls(d.get("numericalFeatures"), d.get("discreteFeatures"), d.get("categoricalFeatures"),
                   d.get("target"), d.get("operation", "training"), d.get("id"))

    @classmethod
    def from_json(cls, s: str | bytes) -> "DataInstance":
        return cls.from_dict(json.loads(s))

    def to_dict(self) -> dict:
        d = {"numericalFeatures": self.numericalFeatures, "discreteFeatures": self.discreteFeatures,
             "categoricalFeatures": self.categoricalFeatures, "target": self.target,
             "operation": self.operation}
        if self.id is not None:
            d["id"] = self.id
        return _clean({k: v for k, v in d.items() if v is not None})

    def to_json(self) -> str:
        return json.dumps(self.to_dict())

    def set_metadata(self, topic, partition, key, offset, timestamp) -> None:
        self.metadata = {"topic": topic, "partition": partition, "key": key, "offset": offset,
                         "timestamp": timestamp}


@dataclass
class QueryResponse:
    """10-argument constructor order fixed by FlinkNetwork.scala:196-208."""
    responseId: int = -1
    bucketIndex: Optional[int] = None
    mlpId: int = -1
    preprocessors: Optional[List[PreprocessorPOJO]] = None
    learner: Optional[LearnerPOJO] = None
    protocol: Optional[str] = None
    dataFitted: Optional[int] = None
    loss: Optional[float] = None
    cumulativeLoss: Optional[float] = None
    score: Optional[float] = None

    def to_dict(self) -> dict:
        return _clean({
            "responseId": self.responseId, "bucketIndex": self.bucketIndex, "mlpId": self.mlpId,
            "preprocessors": [p.to_dict() for p in self.preprocessors] if self.preprocessors is not None else None,
            "learner": self.learner.to_dict() if self.learner else None,
            "protocol": self.protocol, "dataFitted": self.dataFitted, "loss": self.loss,
            "cumulativeLoss": self.cumulativeLoss, "score": self.score})

    def to_json(self) -> str:
        return json.dumps(self.to_dict())

    @classmethod
    def from_dict(cls, d: dict) -> "QueryResponse":
        pps = d.get("preprocessors")
        return cls(d.get("responseId", -1), d.get("bucketIndex"), d.get("mlpId", -1),
                   [PreprocessorPOJO.from_dict(p) for p in pps] if pps is not None else None,
                   LearnerPOJO.from_dict(d.get("learner")), d.get("protocol"), d.get("dataFitted"),
                   d.get("loss"), d.get("cumulativeLoss"), d.get("score"))

    def get_size(self) -> int:
        return 8 * 6 + (self.learner.get_size() if self.learner else 0) + _sizeof(self.protocol) + \
            sum(p.get_size() for p in (self.preprocessors or []))

    def __str__(self) -> str:  # what lands on the `responses` topic (Job.scala:89-96)
        return self.to_json()


@dataclass
class Prediction:
    mlpId: int = -1
    dataInstance: Optional[DataInstance] = None
    prediction: Optional[float] = None

    def to_dict(self) -> dict:
        return _clean({"mlpId": self.mlpId,
                       "dataInstance": self.dataInstance.to_dict() if self.dataInstance else None,
                       "prediction": self.prediction})

    def to_json(self) -> str:
        return json.dumps(self.to_dict())

    def get_size(self) -> int:
        return 16 + _sizeof(self.dataInstance.to_dict() if self.dataInstance else None)

    def __str__(self) -> str:
        return self.to_json()


@dataclass
class Statistics:
    """Per-pipeline performance record (FlinkHub.scala:118-127, StatisticsOperator.scala:99-124)."""
    pipeline: int = -1
    protocol: Optional[str] = None
    meanBufferSize: float = 0.0
    models: int = 0          # models shipped
    bytes: int = 0           # bytes shipped
    blocks: int = 0          # message blocks
    fitted: int = 0
    score: float = 0.0
    learningCurve: List[float] = field(default_factory=list)
    lcx: List[int] = field(default_factory=list)

    # the reference's incremental updaters
    def update_stats(self, models: int, nbytes: int, blocks: int) -> None:
        self.models += models
        self.bytes += nbytes
        self.blocks += blocks

    def update_mean_buffer_size(self, v: float) -> None:
        self.meanBufferSize += v

    def update_fitted(self, n: int) -> None:
        self.fitted += n

    def update_score(self, s: float) -> None:
        self.score += s

    def to_dict(self) -> dict:
        return _clean({"pipeline": self.pipeline, "protocol": self.protocol,
                       "meanBufferSize": self.meanBufferSize, "models": self.models,
                       "bytes": self.bytes, "blocks": self.blocks, "fitted": self.fitted,
                       "score": self.score, "learningCurve": self.learningCurve, "lcx": self.lcx})

    def get_size(self) -> int:
        return 8 * 7 + _sizeof(self.protocol) + 8 * (len(self.learningCurve) + len(self.lcx))


@dataclass
class JobStatistics:
    jobName: str = ""
    parallelism: int = 1
    durationMs: float = 0.0
    statistics: List[Statistics] = field(default_factory=list)
    extra: Dict[str, Any] = field(default_factory=dict)   # records/s, sync GB/s, roofline fraction

    def to_dict(self) -> dict:
        d = {"jobName": self.jobName, "parallelism": self.parallelism, "durationMs": self.durationMs,
             "statistics": [s.to_dict() for s in self.statistics]}
        d.update(self.extra)
        return _clean(d)

    def to_json(self) -> str:
        return json.dumps(self.to_dict())

    def __str__(self) -> str:  # PerformanceWriter.scala:6-8
        return self.to_json()
