"""Streaming service: the reference's topic-in / topic-out contract around an engine.

The reference job is driven entirely through Kafka topics (Job.scala:42-105,127-142): JSON records
on `trainingData` / `forecastingData`, JSON requests on `requests`; answers on `responses`,
`predictions` and `performance`.  `OnlineLearningService` keeps exactly that contract with
in-process topics: producers `put()` JSON strings, consumers `get()` them.  One background thread
plays the Flink job: it drains the input topics, batches records (native JSON scanner ->
staged rows), feeds the engine, and implements the performance harness of
utils/statistics/StatisticsOperator.scala - when `--test true` and the input has been quiet for
`--timeout` ms it runs the termination probe, publishes `JobStatistics` on `performance` and stops
(the reference's JobTerminator, utils/JobTerminator.scala:7-9).

`engine="gpu"` uses `omldm_b200.engine.Engine` (CUDA when available, otherwise its CPU tensor path);
`engine="host"` uses the fp64 message-driven `HostJob` with `parallelism` in-process workers.
"""
from __future__ import annotations

import collections
import json
import threading
import time
from typing import Dict, List, Optional


from .utils.config import TOPICS, JobParameters


class MessageTopic:
    """Unbounded FIFO of strings with blocking reads (one Kafka topic, one partition).  `put_many` / `drain` move whole
    chunks under one lock acquisition - the per-record cost of a Python queue would otherwise bound the stream."""

    def __init__(self, name: str):
        self.name = name
        self._d: "collections.deque[str]" = collections.deque()
        self._cv = threading.Condition()
        self.produced = 0

    def put(self, msg) -> None:
        m = msg if isinstance(msg, str) else json.dumps(msg)
        with self._cv:
            self._d.append(m)
            self.produced += 1
            self._cv.notify()

    def put_many(self, msgs) -> None:
        ms = [m if isinstance(m, str) else json.dumps(m) for m in msgs]
        with self._cv:
            self._d.extend(ms)
            self.produced += len(ms)
            self._cv.notify_all()

    def get(self, timeout: Optional[float] = None) -> Optional[str]:
        with self._cv:
            if not self._d and timeout:
                self._cv.wait_for(lambda: bool(self._d), timeout)
            return self._d.popleft() if self._d else None

    def drain(self, limit: int = 1 << 30) -> List[str]:
        with self._cv:
            n = min(limit, len(self._d))
            if n == len(self._d):
                out = list(self._d)
                self._d.clear()
            else:
                out = [self._d.popleft() for _ in range(n)]
        return out

    def __len__(self) -> int:
        return len(self._d)


class OnlineLearningService:
    def __init__(self, params: Optional[JobParameters] = None, engine: str = "gpu", device=None,
                 batch: int = 4096, poll_ms: float = 2.0):
        self.params = params or JobParameters(parallelism=1)
        self.topics: Dict[str, MessageTopic] = {t: MessageTopic(self.params.topics.get(t, t)) for t in TOPICS}
        self.batch, self.poll_s = batch, poll_ms / 1000.0
        self.kind = engine
        if engine == "host":
            from .runtime.job import HostJob
            self.job = HostJob(self.params)
            self.engine = None
        else:
            from .engine import Engine
            self.job = None
            self.engine = Engine(self.params, 0, 1, device)
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self.terminated = threading.Event()
        self._last_activity: Optional[float] = None
        self._d: Optional[int] = None
        self.error: Optional[BaseException] = None

    # ------------------------------------------------------------------ topic access
    def topic(self, name: str) -> MessageTopic:
        return self.topics[name]

    def start(self) -> "OnlineLearningService":
        self._thread = threading.Thread(target=self._run, name="omldm-job", daemon=True)
        self._thread.start()
        return self

    def stop(self) -> None:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=30)
        if self.engine is not None:
            self.engine.close()

    def wait_terminated(self, timeout: float = 60.0) -> bool:
        return self.terminated.wait(timeout)

    # ------------------------------------------------------------------ the "job"
    def _run(self) -> None:
        try:
            while not self._stop.is_set():
                busy = self._pump_requests()
                busy |= self._pump_training()
                busy |= self._pump_forecasting()
                now = time.monotonic()
                if busy:
                    self._last_activity = now
                elif (self.params.test and self._last_activity is not None
                      and (now - self._last_activity) * 1000.0 >= self.params.timeout):
                    self._terminate()
                    return
                if not busy:
                    time.sleep(self.poll_s)
        except BaseException as e:   # surfaced to the caller instead of dying silently in a thread
            self.error = e
            self.terminated.set()

    def _publish(self, topic: str, msgs) -> None:
        for m in msgs:
            self.topics[topic].put(m)

    def _pump_requests(self) -> bool:
        reqs = self.topics["requests"].drain()
        for r in reqs:
            if self.job is not None:
                n0 = len(self.job.responses)
                self.job.submit_request(r)
                self._publish("responses", self.job.responses[n0:])
            else:
                self._publish("responses", self.engine.handle_request(r))
        return bool(reqs)

    def _rows(self, lines: List[str]):
        from .stream.ring import parse_json_records
        if self._d is None:
            for l in lines:
                try:
                    d = json.loads(l)
                    self._d = len(d.get("numericalFeatures") or []) + len(d.get("discreteFeatures") or [])
                    break
                except Exception:
                    continue
            if self._d is None:
                return None
        text = ("\n".join(lines) + "\n").encode()
        rows, nf, _ = parse_json_records(text, self._d, len(lines))
        return rows

    def _pump_training(self) -> bool:
        lines = self.topics["trainingData"].drain(self.batch)
        if not lines:
            return False
        if self.job is not None:
            for l in lines:
                self.job.submit_data(l)
            return True
        rows = self._rows(lines)
        if rows is not None and rows.shape[0]:
            rows = rows[rows[:, self._d] == 1.0]
            if rows.shape[0]:
                self.engine.train_batch(rows.to(self.engine.device), self._d)
        return True

    def _pump_forecasting(self) -> bool:
        lines = self.topics["forecastingData"].drain(self.batch)
        if not lines:
            return False
        if self.job is not None:
            n0 = len(self.job.predictions)
            for l in lines:
                self.job.submit_data(l)
            self._publish("predictions", self.job.predictions[n0:])
            return True
        from .api.pojos import DataInstance
        good = [l for l in lines if l.lstrip().startswith("{")]
        rows = self._rows(good)
        if rows is not None and rows.shape[0] and self.engine.d is not None:
            rows[:, self._d] = 1.0
            insts = []
            for l in good:
                try:
                    insts.append(DataInstance.from_json(l))
                except Exception:
                    pass
            self._publish("predictions", self.engine.forecast_batch(rows.to(self.engine.device), insts[: rows.shape[0]]))
        return True

    def _terminate(self) -> None:
        if self.job is not None:
            js = self.job.end_of_stream()
        else:
            js = self.engine.terminate()
        if js is not None:
            self.topics["performance"].put(str(js))
        self.terminated.set()
