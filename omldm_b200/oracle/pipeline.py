"""Oracle ML pipeline = ordered preprocessors + one learner, with the reference's holdout rule.

Holdout (operators/spoke/FlinkSpoke.scala:92-107): of every 10 training points, numbers 8 and 9
are appended to a bounded test set; if that evicts an old point the evicted point is trained on.
In the reference the test set belongs to the spoke and is shared by its pipelines; the
`HoldoutSampler` below is therefore separate from `OraclePipeline`.
"""
from __future__ import annotations

from collections import deque
from typing import Deque, List, Optional, Tuple

import numpy as np

from ..api.pojos import LearnerPOJO, PreprocessorPOJO, QueryResponse, Request
from .learners import OracleLearner, make_oracle_learner
from .preprocessors import OraclePreprocessor, make_oracle_preprocessor


class HoldoutSampler:
    def __init__(self, capacity: int = 256):
        self.capacity = capacity
        self.count = 0
        self.test_set: Deque[Tuple[np.ndarray, float]] = deque()

    def route(self, x: np.ndarray, y: float) -> Optional[Tuple[np.ndarray, float]]:
        """Returns the point to train on now (possibly an evicted one) or None."""
        c = self.count % 10
        self.count += 1
        if c >= 8:
            self.test_set.append((x, y))
            if len(self.test_set) > self.capacity:
                return self.test_set.popleft()
            return None
        return (x, y)

    def arrays(self):
        if not self.test_set:
            return np.zeros((0, 0)), np.zeros(0)
        return np.stack([p[0] for p in self.test_set]), np.array([p[1] for p in self.test_set])


class OraclePipeline:
    def __init__(self, request: Request):
        self.request = request
        self.preprocessors: List[OraclePreprocessor] = [
            make_oracle_preprocessor(p.name, p.hyperParameters) for p in (request.preProcessors or [])]
        self.learner: OracleLearner = make_oracle_learner(request.learner.name, request.learner.hyperParameters)
        if request.learner.parameters and request.learner.parameters.get("a") is not None:
            self.learner.from_parameters(request.learner.parameters)
        self.window_loss = 0.0
        self.window_n = 0

    def _pre(self, x: np.ndarray, fit: bool) -> np.ndarray:
        for p in self.preprocessors:
            x = p.fit_transform(x) if fit else p.transform(x)
        return x

    def fit(self, x: np.ndarray, y: float) -> float:
        loss = self.learner.fit(self._pre(np.asarray(x, dtype=np.float64), True), y)
        self.window_loss += loss
        self.window_n += 1
        return loss

    def fit_batch(self, X: np.ndarray, y: np.ndarray) -> float:
        X = np.asarray(X, dtype=np.float64)
        for p in self.preprocessors:
            X = p.fit_transform_batch(X)
        loss = self.learner.fit_batch(X, np.asarray(y, dtype=np.float64))
        self.window_loss += loss
        self.window_n += len(X)
        return loss

    def predict(self, x: np.ndarray) -> float:
        xx = self._pre(np.asarray(x, dtype=np.float64), False)
        self.learner.ensure_init(len(xx))
        return self.learner.predict(xx)

    def score(self, X: np.ndarray, y: np.ndarray) -> float:
        if len(X) == 0:
            return 0.0
        X = np.asarray(X, dtype=np.float64)
        for p in self.preprocessors:
            X = p.transform_batch(X)
        return self.learner.score(X, y)

    def query_response(self, response_id: int, protocol: str, test: Tuple[np.ndarray, np.ndarray]) -> QueryResponse:
        lrn = self.learner
        pojo = LearnerPOJO(lrn.name, lrn.hyper_parameters(),
                           lrn.to_parameters() if lrn.d is not None else None,
                           lrn.data_structure() if hasattr(lrn, "data_structure") and lrn.d is not None else None)
        pps = [PreprocessorPOJO(p.name, dict(p.hyper), p.to_parameters(), None) for p in self.preprocessors]
        loss = self.window_loss / self.window_n if self.window_n else 0.0
        return QueryResponse(response_id, None, self.request.id, pps, pojo, protocol, lrn.fitted, loss,
                             lrn.cum_loss, self.score(*test) if lrn.d is not None else 0.0)
