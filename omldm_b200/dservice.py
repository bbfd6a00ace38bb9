"""Multi-rank streaming service: the reference's topic-in / topic-out contract on N workers.

In the reference, requests arrive at any time on the `requests` topic, are validated once by a single
`PipelineMap` instance and broadcast to every worker while data keeps flowing (Job.scala:127-153,
utils/parsers/requestStream/PipelineMap.scala:31-57, operators/spoke/FlinkSpoke.scala:117-175); every worker
reads its own partitions of `trainingData` / `forecastingData` (Job.scala:42-70) and serves forecasts from its local
replica between training records (FlinkSpoke.scala:105).  `DistributedService` keeps that behaviour on the SPMD
GPU engine - one instance per rank, `torch.distributed` (NCCL or gloo) initialised:

  * rank 0 owns the `requests` topic.  At every step boundary it drains the topic and broadcasts the requests; every
    rank applies them in the same order (Create / Update / Query / Delete in mid-stream), rank 0 publishes the
    responses.
  * every rank owns its partition of `trainingData` / `forecastingData`.  The step schedule is agreed per round: the
    ranks exchange how many records they hold, and if anybody has data EVERY rank trains one batch of the same
    shape (short partitions are padded with empty rows), which is what the barrier protocols need.  Records that
    arrive before the first Create wait in the engine's record buffer.
  * forecasts are answered between training batches from the local replica and published on the rank's
    `predictions` topic.
  * test mode: when no rank has seen input for `timeout` ms the ranks run the termination probe together and rank 0
    publishes `JobStatistics` on `performance` (utils/statistics/StatisticsOperator.scala:93-142).

One control round costs one object broadcast and one small all-reduce; the data plane between rounds is the engine's.
"""
from __future__ import annotations

import json
import time
from typing import Dict, List, Optional

import torch

from .engine import Engine
from .service import MessageTopic
from .utils.config import TOPICS, JobParameters


def _bucket(n: int, lo: int = 64) -> int:
    b = lo
    while b < n:
        b *= 2
    return b


class DistributedService:
    def __init__(self, params: Optional[JobParameters], rank: int, world: int, device, *, batch: int = 4096,
                 impl: str = "fused", group=None, n_features: Optional[int] = None):
        self.params = params or JobParameters(parallelism=world)
        self.rank, self.world, self.group = rank, world, group
        self.engine = Engine(self.params, rank, world, device, impl=impl, group=group)
        self.device = self.engine.device
        self.topics: Dict[str, MessageTopic] = {t: MessageTopic(self.params.topics.get(t, t)) for t in TOPICS}
        self.batch = batch
        self.d = n_features
        self.rounds = 0
        self.steps = 0
        self.terminated = False
        self._last_activity: Optional[float] = None
        self._stop_requested = False

    def topic(self, name: str) -> MessageTopic:
        return self.topics[name]

    def request_stop(self) -> None:
        """Rank 0: end the job at the next round even if test mode is off."""
        self._stop_requested = True

    # ------------------------------------------------------------------ helpers
    def _dist(self):
        import torch.distributed as dist
        return dist if (self.world > 1 and dist.is_available() and dist.is_initialized()) else None

    def _bcast(self, obj):
        d = self._dist()
        if d is None:
            return obj
        box = [obj if self.rank == 0 else None]
        d.broadcast_object_list(box, src=0, group=self.group)
        return box[0]

    def _allmax(self, vals: List[int]) -> List[int]:
        d = self._dist()
        if d is None:
            return vals
        dev = self.device if self.device.type == "cuda" else torch.device("cpu")
        t = torch.tensor(vals, dtype=torch.int64, device=dev)
        d.all_reduce(t, op=d.ReduceOp.MAX, group=self.group)
        return [int(v) for v in t.tolist()]

    def _rows(self, lines: List[str]):
        from .stream.ring import parse_json_records
        if not lines or self.d is None:
            return None
        text = ("\n".join(lines) + "\n").encode()
        rows, _, _ = parse_json_records(text, self.d, len(lines))
        return rows

    def _sniff_dim(self) -> int:
        """Feature count of the first parsable record this rank holds (0 = nothing yet)."""
        if self.d is not None:
            return self.d
        for name in ("trainingData", "forecastingData"):
            q = self.topics[name]
            with q._cv:
                head = list(q._d)[:8]
            for l in head:
                try:
                    o = json.loads(l)
                    return len(o.get("numericalFeatures") or []) + len(o.get("discreteFeatures") or [])
                except Exception:
                    continue
        return 0

    # ------------------------------------------------------------------ one control round
    def step(self) -> bool:
        """One round: requests, at most one training batch per rank, forecasts, termination check.  Collective:
        every rank calls it the same number of times.  Returns False once the job has terminated."""
        if self.terminated:
            return False
        self.rounds += 1
        now = time.monotonic()
        # ---- (1) control plane: rank 0's requests, in arrival order, to everybody
        reqs = self.topics["requests"].drain() if self.rank == 0 else None
        reqs, stop = self._bcast((reqs, self._stop_requested) if self.rank == 0 else None)
        busy = bool(reqs)
        for r in reqs:
            out = self.engine.handle_request(r)
            if self.rank == 0:
                for m in out:
                    self.topics["responses"].put(m)
        # ---- (2) agree on the step: does anybody hold training records, and of which width
        have = min(len(self.topics["trainingData"]), self.batch)
        have_f = len(self.topics["forecastingData"])
        g_have, g_d, g_busy_f = self._allmax([have, self._sniff_dim(), 1 if have_f else 0])
        if self.d is None and g_d > 0:
            self.d = g_d
        if g_have > 0 and self.d:
            lines = [l for l in self.topics["trainingData"].drain(self.batch) if l.lstrip().startswith("{")]
            rows = self._rows(lines)
            B = _bucket(g_have)
            Z = torch.zeros(B, (self.d + 2 + 31) // 32 * 32)
            if rows is not None and rows.shape[0]:
                rows = rows[rows[:, self.d] == 1.0]
                Z[: rows.shape[0]] = rows[:B]
            self.engine.train_batch(Z.to(self.device), self.d)        # same shape on every rank; short ones padded
            self.steps += 1
            busy = True
        # ---- (3) forecasts from the local replica, between training batches
        if have_f and self.d and self.engine.pipelines:
            from .api.pojos import DataInstance
            good = [l for l in self.topics["forecastingData"].drain(self.batch) if l.lstrip().startswith("{")]
            rows = self._rows(good)
            if rows is not None and rows.shape[0]:
                rows[:, self.d] = 1.0
                insts = []
                for l in good:
                    try:
                        insts.append(DataInstance.from_json(l))
                    except Exception:
                        pass
                for m in self.engine.forecast_batch(rows.to(self.device), insts[: rows.shape[0]]):
                    self.topics["predictions"].put(m)
            busy = True
        elif have_f and not self.engine.pipelines:
            pass                                                      # no model yet: forecasts wait in the topic
        # ---- (4) quiescence / termination, decided together
        if busy or g_busy_f and self.engine.pipelines:
            self._last_activity = now
        idle_ms = 0 if self._last_activity is None else int((now - self._last_activity) * 1000.0)
        quiet = 1 if (self.params.test and self._last_activity is not None and idle_ms >= self.params.timeout) else 0
        # every rank must be quiet (MAX of "not quiet" = 0); rank 0's stop request ends the job as well
        (not_quiet,) = self._allmax([0 if quiet else 1])
        if stop or not_quiet == 0:
            js = self.engine.terminate()
            if js is not None and self.rank == 0:
                self.topics["performance"].put(str(js))
            self.terminated = True
            return False
        return True

    def run(self, max_rounds: int = 1 << 62, poll_ms: float = 1.0) -> None:
        """Round after round until termination (test mode quiescence or `request_stop` on rank 0)."""
        n = 0
        while n < max_rounds and self.step():
            n += 1
            if self._last_activity is None or (time.monotonic() - self._last_activity) > 2 * poll_ms / 1000.0:
                time.sleep(poll_ms / 1000.0)

    def close(self) -> None:
        self.engine.close()
