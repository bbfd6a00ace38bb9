"""Peer liveness watchdog ("a per-rank heartbeat word in symmetric memory").

The reference has no failure detector of its own and leaves recovery to Flink.  Here every rank periodically publishes
a counter into all peers' signal pads (`PipelineComm.heartbeat`, one relaxed store per peer from a tiny kernel on a
side stream, never waits) and looks at its own pad: a peer whose counter stopped advancing while this rank's own
beats went on is reported as stalled.  Together with the bounded spins of the protocol kernels (status bits instead
of hangs) this is the failure-detection story of the GPU engine."""
from __future__ import annotations

import threading
import time
from typing import Callable, Dict, List, Optional


class HeartbeatMonitor:
    """Pure bookkeeping: feed it the heartbeat vector from time to time, ask who is stalled."""

    def __init__(self, world: int, rank: int, stall_after_s: float = 5.0, clock: Callable[[], float] = time.monotonic):
        self.world, self.rank, self.stall_after_s, self.clock = world, rank, stall_after_s, clock
        now = clock()
        self.last_value: List[int] = [0] * world
        self.last_change: List[float] = [now] * world

    def observe(self, beats: List[int]) -> None:
        now = self.clock()
        for r, v in enumerate(beats[: self.world]):
            if v != self.last_value[r]:
                self.last_value[r], self.last_change[r] = v, now

    def stalled(self) -> List[int]:
        now = self.clock()
        return [r for r in range(self.world)
                if r != self.rank and now - self.last_change[r] > self.stall_after_s]


class Watchdog:
    """Background thread: beat, observe, call `on_stall(ranks)` when the set of stalled peers changes."""

    def __init__(self, comm, rank: int, period_s: float = 0.5, stall_after_s: float = 5.0,
                 on_stall: Optional[Callable[[List[int]], None]] = None):
        self.comm, self.period_s, self.on_stall = comm, period_s, on_stall
        self.monitor = HeartbeatMonitor(comm.k, rank, stall_after_s)
        self.beat = 0
        self.reported: List[int] = []
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._stream = None

    def tick(self) -> List[int]:
        """One beat + observation (also usable without the thread)."""
        import torch
        self.beat += 1
        if self.comm.device.type == "cuda":
            if self._stream is None:
                self._stream = torch.cuda.Stream(device=self.comm.device)
            with torch.cuda.stream(self._stream):
                self.comm.heartbeat(self.beat)
                self._stream.synchronize()
        else:
            self.comm.heartbeat(self.beat)
        self.monitor.observe(self.comm.heartbeats())
        bad = self.monitor.stalled()
        if bad != self.reported:
            self.reported = bad
            if self.on_stall is not None:
                self.on_stall(bad)
        return bad

    def start(self) -> "Watchdog":
        def loop():
            while not self._stop.wait(self.period_s):
                self.tick()
        self._thread = threading.Thread(target=loop, name="omldm-watchdog", daemon=True)
        self._thread.start()
        return self

    def stop(self) -> None:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=5.0)

    def snapshot(self) -> Dict[str, object]:
        return {"beat": self.beat, "seen": list(self.monitor.last_value), "stalled": list(self.reported)}
