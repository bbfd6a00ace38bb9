"""Checkpoint / resume.

Reference: utils/Checkpointing.scala:11-23 enables Flink checkpointing (`--checkpointing true`,
`--checkInterval` ms, `--stateBackend` path); the spoke snapshots its pipelines, test set and
buffers (FlinkSpoke.scala:233-251), the hub its parameter-server node (FlinkHub.scala:36-52), the
registry its pipeline map (PipelineMap.scala:59-64).  Here every rank writes one file with
{pipeline specs, learner tensors, preprocessor statistics, holdout ring, record count}, rank 0 adds
a manifest; resume = re-Create from the specs, then load the tensors.  The restore bug of the
reference (restored pipelines dropped, FlinkSpoke.scala:289-305) is not reproduced.
"""
from __future__ import annotations

import json
import os
import time
from typing import Optional

import torch


def _rank_file(directory: str, rank: int) -> str:
    return os.path.join(directory, f"rank{rank:03d}.pt")


def save_checkpoint(engine, directory: str, step: int = 0) -> str:
    os.makedirs(directory, exist_ok=True)
    if engine.device.type == "cuda":
        torch.cuda.synchronize(engine.device)
    path = _rank_file(directory, engine.rank)
    tmp = path + ".tmp"
    torch.save({"step": step, "time": time.time(), "engine": engine.state_dict()}, tmp)
    os.replace(tmp, path)                      # atomic: a crash never leaves a torn checkpoint
    if engine.rank == 0:
        man = {"world": engine.world, "step": step, "time": time.time(), "job": engine.params.jobName,
               "pipelines": sorted(engine.pipelines)}
        with open(os.path.join(directory, "manifest.json.tmp"), "w") as f:
            json.dump(man, f)
        os.replace(os.path.join(directory, "manifest.json.tmp"), os.path.join(directory, "manifest.json"))
    return path


def load_checkpoint(engine, directory: str) -> Optional[int]:
    """Restore this rank's state; returns the step recorded in the checkpoint (None if absent).
    A job restarted with fewer ranks restores rank r from file r (model replicas are identical after
    a synchronisation, so dropping the extra files loses only their un-synced deltas)."""
    path = _rank_file(directory, engine.rank)
    if not os.path.exists(path):
        path = _rank_file(directory, 0)
        if not os.path.exists(path):
            return None
    blob = torch.load(path, map_location="cpu", weights_only=False)
    engine.load_state_dict(blob["engine"])
    return int(blob.get("step", 0))


class PeriodicCheckpointer:
    """`--checkInterval` milliseconds between snapshots (Checkpointing.scala:15)."""

    def __init__(self, engine, directory: str, interval_ms: int = 5000):
        self.engine, self.directory, self.interval = engine, directory, interval_ms / 1000.0
        self.last = time.monotonic()
        self.saved = 0

    def maybe_save(self, step: int) -> bool:
        now = time.monotonic()
        if now - self.last < self.interval:
            return False
        save_checkpoint(self.engine, self.directory, step)
        self.last = now
        self.saved += 1
        return True
