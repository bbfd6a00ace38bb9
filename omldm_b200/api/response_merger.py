"""Merge the per-worker answers to one Query into one response.

Reference: utils/ResponseConstructor.scala:17-61 - group by responseId; once `parallelism`
answers arrived: SUM dataFitted, MEAN of loss / cumulativeLoss / score, any non-null
preprocessors / learner / protocol.  Id -1 (termination probe) and bucket fragments
(`loss is None`) pass through unmerged (:20-21).
"""
from __future__ import annotations

from typing import Dict, List, Optional

from .pojos import QueryResponse


class ResponseConstructor:
    def __init__(self, parallelism: int):
        self.parallelism = parallelism
        self.pending: Dict[int, List[QueryResponse]] = {}
        self.expected: Dict[int, int] = {}

    def expect(self, response_id: int, n: int) -> None:
        """Worker-0-only queries (HT / K-means) are answered by a single worker."""
        self.expected[response_id] = n

    def flat_map(self, resp: QueryResponse) -> List[QueryResponse]:
        if resp.responseId == -1 or resp.loss is None:
            return [resp]
        lst = self.pending.setdefault(resp.responseId, [])
        lst.append(resp)
        need = self.expected.get(resp.responseId, self.parallelism)
        if len(lst) < need:
            return []
        del self.pending[resp.responseId]
        self.expected.pop(resp.responseId, None)
        return [self.merge(lst)]

    @staticmethod
    def merge(lst: List[QueryResponse]) -> QueryResponse:
        n = len(lst)

        def first(attr):
            return next((getattr(r, attr) for r in lst if getattr(r, attr) is not None), None)

        def mean(attr) -> Optional[float]:
            vals = [getattr(r, attr) for r in lst if getattr(r, attr) is not None]
            return sum(vals) / n if vals else None

        return QueryResponse(
            responseId=lst[0].responseId, bucketIndex=lst[0].bucketIndex, mlpId=lst[0].mlpId,   # head's index kept (:35)
            preprocessors=first("preprocessors"), learner=first("learner"), protocol=first("protocol"),
            dataFitted=sum(int(r.dataFitted or 0) for r in lst),
            loss=mean("loss"), cumulativeLoss=mean("cumulativeLoss"), score=mean("score"))
