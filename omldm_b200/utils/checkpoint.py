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


def _barrier(engine) -> None:
    if engine.world > 1:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            dist.barrier(group=engine.group)


def save_checkpoint(engine, directory: str, step: int = 0) -> str:
    """One consistent snapshot.  COLLECTIVE when world > 1: every rank calls it at the SAME step (the reference's
    checkpoints are barrier aligned).  The protocols are settled first (`engine.flush()`), the ranks meet, every rank
    writes its file, they meet again, and only then does rank 0 write the manifest - a manifest therefore always
    describes a complete set of rank files of one step.  `step` is the stream position to resume from."""
    os.makedirs(directory, exist_ok=True)
    engine.flush()
    if engine.device.type == "cuda":
        torch.cuda.synchronize(engine.device)
    _barrier(engine)
    path = _rank_file(directory, engine.rank)
    tmp = path + ".tmp"
    torch.save({"step": step, "time": time.time(), "engine": engine.state_dict()}, tmp)
    os.replace(tmp, path)                      # atomic: a crash never leaves a torn checkpoint
    _barrier(engine)
    if engine.rank == 0:
        man = {"world": engine.world, "step": step, "time": time.time(), "job": engine.params.jobName,
               "pipelines": sorted(engine.pipelines)}
        with open(os.path.join(directory, "manifest.json.tmp"), "w") as f:
            json.dump(man, f)
        os.replace(os.path.join(directory, "manifest.json.tmp"), os.path.join(directory, "manifest.json"))
    return path


_ADDITIVE = ("ORR", "K-means")     # learners whose `buf` is a delta of additive statistics (merge_mode "sum")


def _merge_holdout(rings, counts, cap: int):
    """Interleaving merge of the valid rows of several holdout rings, capped at `cap` rows
    (SpokeLogic.scala:37-50 / DataSet.merge).  Returns (ring, rows_valid)."""
    from ..ops.functional import holdouts_before
    parts = []
    for ring, c in zip(rings, counts):
        if ring is not None:
            n = min(holdouts_before(int(c)), ring.shape[0])
            if n > 0:
                parts.append(ring[:n])
    if not parts:
        return (None if all(r is None for r in rings) else next(r for r in rings if r is not None).clone()), 0
    rows, i = [], 0
    while len(rows) < cap and any(i < p.shape[0] for p in parts):
        for p in parts:
            if i < p.shape[0] and len(rows) < cap:
                rows.append(p[i])
        i += 1
    out = torch.zeros(cap, parts[0].shape[1], dtype=parts[0].dtype)
    m = len(rows)
    out[:m] = torch.stack(rows)
    return out, m


def _merge_scaler(states):
    """Chan et al. pairwise merge of (count, mean, M2) running moments."""
    acc = states[0].clone().double()
    d = (acc.numel() - 1) // 2
    for s in states[1:]:
        s = s.double()
        na, nb = float(acc[0]), float(s[0])
        if nb == 0:
            continue
        if na == 0:
            acc = s.clone()
            continue
        n = na + nb
        delta = s[1:1 + d] - acc[1:1 + d]
        acc[1 + d:] = acc[1 + d:] + s[1 + d:] + delta * delta * (na * nb / n)
        acc[1:1 + d] = acc[1:1 + d] + delta * (nb / n)
        acc[0] = n
    return acc.to(states[0].dtype)


def _merge_pipeline(pss):
    """Merge the saved replicas of ONE pipeline (the reference's `node.merge(remoteStates)`,
    FlinkSpoke.scala:291-305): averaged models are averaged, additive statistics keep the global part
    and sum the un-synchronised deltas, counters add up, scaler moments merge exactly."""
    built = [ps for ps in pss if ps.get("learner") is not None]
    if not built:
        return dict(pss[0])
    out = dict(built[0])
    name = out["request"]["learner"]["name"]
    ls = [ps["learner"] for ps in built]
    l = dict(ls[0])
    if name != "HT" and l.get("buf") is not None and len(ls) > 1:
        bufs = torch.stack([x["buf"].double() for x in ls])
        l["buf"] = (bufs.sum(0) if name in _ADDITIVE else bufs.mean(0)).to(ls[0]["buf"].dtype)
    l["stats"] = torch.stack([x["stats"].double() for x in ls]).sum(0).to(ls[0]["stats"].dtype)
    out["learner"] = l
    out["batches"] = sum(int(ps.get("batches", 0)) for ps in built)
    pre = []
    for i, p0 in enumerate(out.get("pre") or []):
        names = [pp["name"] for pp in out["request"].get("preProcessors") or []]
        nm = names[i] if i < len(names) else ""
        sts = [ps["pre"][i]["state"] for ps in built if ps.get("pre") and "state" in ps["pre"][i]]
        if nm == "StandardScaler" and sts:
            pre.append({"state": _merge_scaler(sts)})
        elif nm == "MinMaxScaler" and sts:
            d = sts[0].numel() // 2
            st = torch.stack(sts)
            pre.append({"state": torch.cat([st[:, :d].min(0).values, st[:, d:].max(0).values])})
        else:
            pre.append(p0)
    out["pre"] = pre if out.get("pre") is not None else None
    return out


def merge_engine_states(states):
    """Fold the states of several old workers into one (restore with fewer ranks than were saved)."""
    if len(states) == 1:
        return states[0]
    first = states[0]
    cap = next((s["holdout"].shape[0] for s in states if s["holdout"] is not None), 0)
    ring, m = _merge_holdout([s["holdout"] for s in states], [s["count"] for s in states], cap) if cap else (None, 0)
    from ..ops.functional import holdouts_before
    count = sum(int(s["count"]) for s in states)
    if ring is not None:
        if m < cap:
            count = min(count, 5 * m + 10)
        while count > 0 and min(holdouts_before(count), cap) > m:   # never expose unfilled ring rows as valid
            count -= 1
    out = {"rank": first["rank"], "world": first["world"], "count": count, "d": first["d"],
           "sparse_nnz": first.get("sparse_nnz"),
           "registry": first["registry"], "holdout": ring, "pipelines": {}}
    for pid in first["pipelines"]:
        out["pipelines"][pid] = _merge_pipeline([s["pipelines"][pid] for s in states if pid in s["pipelines"]])
    return out


def fresh_replica_state(sd):
    """State for a rank that did not exist when the checkpoint was written (restore with MORE ranks):
    it adopts the model but none of the old worker's history - no counters, no pending delta, no
    holdout - so nothing is counted twice."""
    out = dict(sd)
    out["count"], out["holdout"] = 0, (None if sd["holdout"] is None else torch.zeros_like(sd["holdout"]))
    pipes = {}
    for pid, ps in sd["pipelines"].items():
        ps = dict(ps)
        if ps.get("learner") is not None:
            l = dict(ps["learner"])
            l["stats"] = torch.zeros_like(l["stats"])
            if ps["request"]["learner"]["name"] in _ADDITIVE and l.get("buf") is not None:
                l["buf"] = torch.zeros_like(l["buf"])
            ps["learner"], ps["batches"] = l, 0
        pipes[pid] = ps
    out["pipelines"] = pipes
    return out


def load_checkpoint(engine, directory: str) -> Optional[int]:
    """Restore this rank's state; returns the step recorded in the checkpoint (None if absent).

    Rescaling (FlinkSpoke.scala:289-305, SpokeLogic.scala:37-50): restarted with FEWER ranks, rank r
    folds the files r, r + world, r + 2 world, ... into one state (`merge_engine_states`); restarted
    with MORE ranks, a new rank adopts the model of file r mod old_world with empty history
    (`fresh_replica_state`)."""
    old_world = None
    man = os.path.join(directory, "manifest.json")
    if os.path.exists(man):
        with open(man) as f:
            old_world = int(json.load(f).get("world", 0)) or None
    if old_world is None:
        old_world = 0
        while os.path.exists(_rank_file(directory, old_world)):
            old_world += 1
    if old_world == 0:
        return None
    mine = [r for r in range(engine.rank, old_world, engine.world) if os.path.exists(_rank_file(directory, r))]
    if mine:
        blobs = [torch.load(_rank_file(directory, r), map_location="cpu", weights_only=False) for r in mine]
        engine.load_state_dict(merge_engine_states([b["engine"] for b in blobs]))
        return int(blobs[0].get("step", 0))
    src = _rank_file(directory, engine.rank % old_world)
    if not os.path.exists(src):
        return None
    blob = torch.load(src, map_location="cpu", weights_only=False)
    engine.load_state_dict(fresh_replica_state(blob["engine"]))
    return int(blob.get("step", 0))


class PeriodicCheckpointer:
    """`--checkInterval` milliseconds between snapshots (Checkpointing.scala:15).  With several ranks the decision is
    rank 0's: every `agree_every` steps it tells the others whether the interval has elapsed (one tiny broadcast), so
    every rank snapshots at the same step."""

    def __init__(self, engine, directory: str, interval_ms: int = 5000, agree_every: int = 8):
        self.engine, self.directory, self.interval = engine, directory, interval_ms / 1000.0
        self.last = time.monotonic()
        self.saved = 0
        self.agree_every = max(1, agree_every)

    def maybe_save(self, step: int) -> bool:
        eng = self.engine
        if eng.world > 1:
            if step % self.agree_every:
                return False
            import torch.distributed as dist
            dev = eng.device if eng.device.type == "cuda" else torch.device("cpu")
            flag = torch.tensor([1 if (eng.rank == 0 and time.monotonic() - self.last >= self.interval) else 0], device=dev)
            dist.broadcast(flag, 0, group=eng.group)
            due = bool(int(flag.item()))
        else:
            due = time.monotonic() - self.last >= self.interval
        if not due:
            return False
        save_checkpoint(eng, self.directory, step)
        self.last = time.monotonic()
        self.saved += 1
        return True
