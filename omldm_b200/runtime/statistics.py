"""Performance harness: activity tracking, quiescence detection, termination, JobStatistics.

Reference: utils/statistics/StatisticsOperator.scala (activity :74-91, terminate :93-131, timer
:135-142), state/StateAccumulators.scala:54-126 (`StatisticsAggregateFunction`: merge per-hub
statistics - learning curve of hub 0 concatenated, blocks/models/fitted averaged over hubs),
utils/JobTerminator.scala:7-9 (any record on `performance` kills the job),
utils/PerformanceWriter.scala:6-8.
"""
from __future__ import annotations

from typing import Dict, List, Optional

from ..api.pojos import JobStatistics, Statistics


class JobTerminated(Exception):
    """Raised by JobTerminator; the reference's deliberate crash-as-shutdown."""


class StatisticsAggregate:
    def __init__(self):
        self.by_pipeline: Dict[int, Dict[int, Statistics]] = {}

    def add(self, pipeline: int, hub: int, st: Statistics) -> None:
        self.by_pipeline.setdefault(pipeline, {})[hub] = st

    def result(self) -> List[Statistics]:
        out = []
        for pid, hubs in sorted(self.by_pipeline.items()):
            n = len(hubs)
            h0 = hubs.get(0, next(iter(hubs.values())))
            merged = Statistics(pid, h0.protocol)
            merged.models = int(sum(s.models for s in hubs.values()) / n)
            merged.blocks = int(sum(s.blocks for s in hubs.values()) / n)
            merged.fitted = int(sum(s.fitted for s in hubs.values()) / n)
            merged.bytes = int(sum(s.bytes for s in hubs.values()))
            merged.learningCurve = list(h0.learningCurve)
            merged.lcx = list(h0.lcx)
            out.append(merged)
        return out


class StatisticsOperator:
    def __init__(self, job_name: str, parallelism: int, timeout_ms: int, test_set_size: int):
        self.job_name, self.parallelism, self.timeout_ms = job_name, parallelism, timeout_ms
        self.test_set_size = test_set_size
        self.start: Optional[float] = None
        self.end: Optional[float] = None
        self.timer: Optional[float] = None
        self.agg = StatisticsAggregate()
        self.terminate_stats: Dict[int, List[Statistics]] = {}
        self.probe_sent = False

    # activity path (:74-91)
    def on_activity(self, ts_ms: float, pipeline: Optional[int] = None, hub: int = 0,
                    st: Optional[Statistics] = None) -> None:
        if st is not None:
            if self.start is None:
                self.start = ts_ms
            self.agg.add(pipeline, hub, st)
        if self.start is not None:
            self.end = ts_ms
        self.timer = ts_ms + self.timeout_ms

    # timer (:135-142): stream quiet for `timeout` ms -> termination probe
    def on_time(self, now_ms: float) -> bool:
        if self.timer is not None and now_ms >= self.timer and not self.probe_sent:
            self.probe_sent = True
            return True
        return False

    # terminate path (:93-131)
    def on_terminate(self, st: Statistics, n_pipelines: int) -> Optional[JobStatistics]:
        self.terminate_stats.setdefault(st.pipeline, []).append(st)
        got = sum(len(v) for v in self.terminate_stats.values())
        if got < self.parallelism * max(1, n_pipelines):
            return None
        merged = {s.pipeline: s for s in self.agg.result()}
        out = []
        for pid, lst in sorted(self.terminate_stats.items()):
            s = merged.get(pid, Statistics(pid))
            s.meanBufferSize = sum(x.meanBufferSize for x in lst)
            s.fitted = max(s.fitted, sum(x.fitted for x in lst))
            s.score = sum(x.score for x in lst) / max(1, len(lst))
            out.append(s)
        dur = (self.end - self.start) if (self.start is not None and self.end is not None) else 0.0
        return JobStatistics(self.job_name, self.parallelism, dur, out)


def job_terminator(record) -> None:
    raise JobTerminated(str(record))


def performance_writer(js: JobStatistics) -> str:
    return str(js)
