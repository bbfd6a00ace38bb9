"""Query-response bucketing: big parameter maps are cut into <= 10 000-element buckets.

Reference: network/FlinkNetwork.scala:48-149 (`split`: bucket size :50, key format
`name[start-end]` :82,130, scalars unwrapped when a slice has length 1) and :151-240
(`sendQueryResponse`: a response with at most two buckets is sent whole :179,187-191; bucket 0
carries only the learner maps; the bucket whose index equals `#parameter buckets - 1` carries
preprocessors, protocol, dataFitted, loss, cumulativeLoss and score :193-237).

Deliberate fix: when `len % 10000 == 0` the reference computes an empty last slice
(FlinkNetwork.scala:76-81); here the last bucket always ends at `len - 1`.
"""
from __future__ import annotations

import copy
from typing import Any, Dict, List, Optional

from .pojos import LearnerPOJO, QueryResponse

MAX_PARAM_BUCKET_SIZE = 10000


def _as_sequence(v: Any):
    if v is None:
        return []
    if isinstance(v, str):
        return v
    if hasattr(v, "tolist"):
        v = v.tolist()
    if isinstance(v, (list, tuple)):
        return list(v)
    return [v]


def _unwrap(seq):
    if isinstance(seq, str):
        return seq[0] if len(seq) == 1 else seq
    return seq[0] if len(seq) == 1 else seq


def split(params: Optional[Dict[str, Any]], bucket_size: int = MAX_PARAM_BUCKET_SIZE) -> Dict[int, Dict[str, Any]]:
    """bucket index -> {key -> slice}.  Small entries land whole in bucket 0."""
    out: Dict[int, Dict[str, Any]] = {}
    if params is None:
        return out
    for name, value in params.items():
        seq = _as_sequence(value)
        n = len(seq)
        buckets = n // bucket_size + (0 if n % bucket_size == 0 else 1)
        if buckets <= 1:
            out.setdefault(0, {})[name] = _unwrap(seq) if n > 0 else seq
            continue
        for i in range(buckets):
            start = i * bucket_size
            end = min(n, (i + 1) * bucket_size) - 1
            out.setdefault(i, {})[f"{name}[{start}-{end}]"] = _unwrap(seq[start:end + 1])
    return out


def bucket_response(resp: QueryResponse, bucket_size: int = MAX_PARAM_BUCKET_SIZE) -> List[QueryResponse]:
    """The list of QueryResponse fragments that go to the `responses` topic for one answer."""
    if resp.responseId == -1 or resp.learner is None:
        return [resp]
    bp = split(resp.learner.parameters, bucket_size)
    bh = split(resp.learner.hyperParameters, bucket_size)
    bs = split(resp.learner.dataStructure, bucket_size)
    max_buckets = max(len(bp), len(bh), len(bs)) - 1
    if max_buckets < 2:
        return [resp]
    out = []
    last_param_bucket = len(bp) - 1
    for i in range(max_buckets + 1):
        learner = LearnerPOJO(resp.learner.name, copy.copy(bh.get(i)), copy.copy(bp.get(i)), copy.copy(bs.get(i)))
        carries_scalars = (i != 0 and i == last_param_bucket)
        out.append(QueryResponse(
            responseId=resp.responseId, bucketIndex=i, mlpId=resp.mlpId,
            preprocessors=resp.preprocessors if carries_scalars else None,
            learner=learner,
            protocol=resp.protocol if carries_scalars else None,
            dataFitted=resp.dataFitted if carries_scalars else None,
            loss=resp.loss if carries_scalars else None,
            cumulativeLoss=resp.cumulativeLoss if carries_scalars else None,
            score=resp.score if carries_scalars else None))
    return out


def reassemble(fragments: List[QueryResponse]) -> QueryResponse:
    """Inverse of `bucket_response` (client-side helper; not in the reference)."""
    import re
    if len(fragments) == 1 and fragments[0].bucketIndex is None:
        return fragments[0]
    frags = sorted(fragments, key=lambda f: f.bucketIndex or 0)
    pat = re.compile(r"^(.*)\[(\d+)-(\d+)\]$")

    def merge(attr):
        acc: Dict[str, Any] = {}
        parts: Dict[str, list] = {}
        for f in frags:
            m = getattr(f.learner, attr) if f.learner else None
            for k, v in (m or {}).items():
                mm = pat.match(k)
                if mm:
                    parts.setdefault(mm.group(1), []).append((int(mm.group(2)), v if isinstance(v, list) else [v]))
                else:
                    acc[k] = v
        for k, chunks in parts.items():
            acc[k] = [x for _, c in sorted(chunks, key=lambda t: t[0]) for x in c]
        return acc or None

    scal = next((f for f in frags if f.dataFitted is not None or f.protocol is not None), frags[-1])
    return QueryResponse(frags[0].responseId, None, frags[0].mlpId, scal.preprocessors,
                         LearnerPOJO(frags[0].learner.name, merge("hyperParameters"), merge("parameters"),
                                     merge("dataStructure")),
                         scal.protocol, scal.dataFitted, scal.loss, scal.cumulativeLoss, scal.score)
