"""Fake symmetric-memory backend: the fused protocol kernels of csrc/sync.cu restated in Python over
plain host arrays, one thread per rank.

Purpose (SURVEY §4 "protocol tests without GPUs", §7.1 "comm/ ... fake CPU backend"): the device
kernels synchronise through flags in peer memory (monotonic sequence numbers in signal pads, votes,
clocks, an FGM hub block).  Those state machines - not the arithmetic - are where the bugs live, so
they are mirrored here statement by statement and exercised under real thread interleavings on a
CPU-only box.  The pad layout and the algorithms are the ones documented in include/omldm_kernels.h;
whenever sync.cu changes, this file changes with it.

Every `FakeRank` method corresponds to one kernel launch of one rank:

    allreduce(mode, scale)          sync_allreduce_kernel (Synchronous; mode 1 = additive deltas)
    gm(threshold, ...)              sync_allreduce_kernel with the GM vote
    async_push_pull(mode, scale)    sync_async_kernel
    ssp(mode, scale, staleness)     sync_async_kernel + clocks
    easgd(alpha)                    sync_easgd_kernel
    fgm_step(threshold, eps_psi)    sync_fgm_kernel
    barrier() / query_reduce(vals)  device_barrier_kernel / query_reduce_kernel
"""
from __future__ import annotations

import math
import threading
import time
from typing import List, Optional

import numpy as np

PAD_ENTRY, PAD_EXIT, PAD_CLOCK, PAD_VOTE, PAD_PHI, PAD_FGM, PAD_STATUS, PAD_QUERY = 0, 16, 32, 48, 64, 80, 96, 128
PAD_QENTRY, PAD_QEXIT = 256, 272
PAD_WORDS = 512


class FakeWorld:
    """The symmetric arena: per rank a signal pad, a model/delta buffer and a hub buffer."""

    def __init__(self, k: int, n: int, spin_timeout_s: float = 5.0, jitter: float = 0.0, seed: int = 0):
        """`jitter` > 0 makes every rank pause for a random time (up to `jitter` seconds) at the points where
        the kernels touch peer memory, so repeated runs explore different interleavings."""
        self.k, self.n = k, n
        self.jitter = jitter
        self._rngs = [np.random.default_rng(seed * 1000 + r) for r in range(k)]
        self.pads = [np.zeros(PAD_WORDS, dtype=np.float64) for _ in range(k)]   # words hold ints or floats
        self.bufs = [np.zeros(n, dtype=np.float64) for _ in range(k)]
        self.hubs = [np.zeros(n, dtype=np.float64) for _ in range(k)]
        self.lock = threading.Lock()            # stands in for hardware atomics
        self.spin_timeout_s = spin_timeout_s
        self.ranks = [FakeRank(self, r) for r in range(k)]

    def pause(self, r: int) -> None:
        if self.jitter > 0.0:
            g = self._rngs[r]
            if g.random() < 0.5:
                time.sleep(float(g.random()) * self.jitter)

    def shard(self, r: int):
        per = (self.n + self.k - 1) // self.k
        return min(self.n, per * r), min(self.n, per * (r + 1))

    def run(self, fn, *args):
        """Launch `fn(rank, *args)` on every rank concurrently (one 'kernel' per rank)."""
        errs: List[BaseException] = []

        def body(rank):
            try:
                fn(rank, *args)
            except BaseException as e:   # pragma: no cover - reported below
                errs.append(e)
        ts = [threading.Thread(target=body, args=(r,)) for r in self.ranks]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        if errs:
            raise errs[0]


class FakeRank:
    def __init__(self, world: FakeWorld, me: int):
        self.w, self.me = world, me
        self.e = np.zeros(world.n, dtype=np.float64)   # private estimate / global statistics
        self.epoch = 0        # device-managed sequence numbers (the kernels keep them in `local`)
        self.qepoch = 0       # query / termination barrier: its own counter and flag words
        self.clock = 0
        self.step = 0
        self.fgm = {"round": 0, "sub": -1, "zeta": 0.0, "sent": 0, "rounds": 0}

    # ---- flag primitives -------------------------------------------------------------------------
    def _signal_all(self, base: int, seq: int) -> None:
        self.w.pause(self.me)
        for p in range(self.w.k):                      # st.release.sys into every peer's pad
            self.w.pads[p][base + self.me] = seq

    def _wait_all(self, base: int, seq: int) -> None:
        t0 = time.monotonic()
        for t in range(self.w.k):                      # ld.acquire.sys polling of the own pad
            while self.w.pads[self.me][base + t] < seq:
                if time.monotonic() - t0 > self.w.spin_timeout_s:
                    self.w.pads[self.me][PAD_STATUS] = 1
                    return
                time.sleep(0)

    @property
    def buf(self) -> np.ndarray:
        return self.w.bufs[self.me]

    # ---- Synchronous / GM ---------------------------------------------------------------------------
    def allreduce(self, mode: int = 0, scale: Optional[float] = None, gm_threshold: float = -1.0) -> bool:
        w, me = self.w, self.me
        scale = (1.0 if mode == 1 else 1.0 / w.k) if scale is None else scale
        self.epoch += 1
        seq_entry, seq_exit = 2 * self.epoch + 1, 2 * self.epoch + 2
        if gm_threshold >= 0.0:
            d = self.buf - (self.e if mode == 0 else 0.0)
            vote = 2 * self.epoch + (1 if float(d @ d) > gm_threshold else 0)
            for p in range(w.k):
                w.pads[p][PAD_VOTE + me] = vote       # the vote travels before the entry flag
        self._signal_all(PAD_ENTRY, seq_entry)
        self._wait_all(PAD_ENTRY, seq_entry)
        if gm_threshold >= 0.0:
            if not any(w.pads[me][PAD_VOTE + t] == 2 * self.epoch + 1 for t in range(w.k)):
                return False                           # every rank sees the same votes
        lo, hi = w.shard(me)                           # owner-side reduce + write-back of the own shard
        acc = scale * sum(w.bufs[p][lo:hi] for p in range(w.k))
        for p in range(w.k):
            w.bufs[p][lo:hi] = acc
        self._signal_all(PAD_EXIT, seq_exit)
        self._wait_all(PAD_EXIT, seq_exit)
        if mode == 0:                                  # local tail
            self.e[:] = self.buf
        else:
            self.e += self.buf
            self.buf[:] = 0.0
        return True

    def gm(self, threshold: float, mode: int = 0) -> bool:
        return self.allreduce(mode=mode, gm_threshold=threshold)

    # ---- Asynchronous / SSP / EASGD --------------------------------------------------------------
    def async_push_pull(self, mode: int = 0, scale: Optional[float] = None) -> None:
        w = self.w
        scale = (1.0 if mode == 1 else 1.0 / w.k) if scale is None else scale
        for owner in range(w.k):
            w.pause(self.me)
            lo, hi = w.shard(owner)
            d = scale * (self.buf[lo:hi] - self.e[lo:hi]) if mode == 0 else scale * self.buf[lo:hi]
            with w.lock:                               # atom.add returning the old value
                old = w.hubs[owner][lo:hi].copy()
                w.hubs[owner][lo:hi] += d
            new = old + d
            self.e[lo:hi] = new
            self.buf[lo:hi] = new if mode == 0 else 0.0

    def ssp(self, staleness: int, mode: int = 0, scale: Optional[float] = None) -> None:
        self.async_push_pull(mode, scale)
        self.clock += 1
        self._signal_all(PAD_CLOCK, self.clock)
        t0 = time.monotonic()
        for t in range(self.w.k):                      # nobody may be more than `staleness` clocks behind me
            while self.w.pads[self.me][PAD_CLOCK + t] + staleness < self.clock:
                if time.monotonic() - t0 > self.w.spin_timeout_s:
                    self.w.pads[self.me][PAD_STATUS] = 2
                    return
                time.sleep(0)

    def easgd(self, alpha: float) -> None:
        w = self.w
        for owner in range(w.k):
            lo, hi = w.shard(owner)
            t = w.hubs[owner][lo:hi].copy()            # peer load of the centre shard
            d = alpha * (self.buf[lo:hi] - t)
            self.buf[lo:hi] -= d
            with w.lock:                               # red.add
                w.hubs[owner][lo:hi] += d

    # ---- FGM ----------------------------------------------------------------------------------------------
    def fgm_init_hub(self, threshold: float) -> None:
        """Host-side initialisation the driver performs on rank 0's pad."""
        pad0 = self.w.pads[0]
        pad0[PAD_FGM + 2] = math.sqrt(threshold) / 2.0
        pad0[PAD_PHI:PAD_PHI + 16] = -math.sqrt(threshold)

    def fgm_step(self, threshold: float, eps_psi: float = 0.01, mode: int = 0, scale: Optional[float] = None,
                 drain: int = 0) -> bool:
        """One monitoring step.  `drain` = g > 0 marks an end-of-stream step of flush generation g: the rank
        tells the hub it has no more data; once the hub has heard that from everybody and no round is open it
        writes g into every pad's terminate word instead of evaluating the counters."""
        w, me, k, st = self.w, self.me, self.w.k, self.fgm
        scale = (1.0 if mode == 1 else 1.0 / k) if scale is None else scale
        hub = w.pads[0]
        phi0 = -math.sqrt(threshold)
        self.step += 1
        d = self.buf - (self.e if mode == 0 else 0.0)
        phi = math.sqrt(float(d @ d)) + phi0
        rnd = st["round"]
        seen_round = int(w.pads[me][PAD_FGM + 3])      # sampled BEFORE this step's hub logic
        # a round announced by the hub in ITS step s is not executed by a rank whose own step number is s: ranks
        # that advance in lock-step (one launch for all emulated ranks) then all execute it in step s+1, whether
        # they sampled the flag before or after the hub wrote it; everybody else acts at first sight
        act = seen_round != rnd and self.step != int(w.pads[me][PAD_FGM + 4])
        w.pause(me)
        if drain:
            hub[PAD_VOTE + me] = drain                 # "no more data from me" (hub's pad only)
        hub[PAD_PHI + me] = phi                        # push-based polling
        sub = int(hub[PAD_FGM + 1])
        theta = hub[PAD_FGM + 2] if hub[PAD_FGM + 2] > 0 else -phi0 / 2.0
        if st["sub"] != sub:                           # new sub-round (or first step): re-anchor
            st["sub"], st["zeta"], st["sent"] = sub, phi, 0
        ci = max(st["sent"], int(math.floor((phi - st["zeta"]) / theta)))
        if ci > st["sent"]:
            with w.lock:
                hub[PAD_FGM + 0] += ci - st["sent"]
            st["sent"] = ci
        w.pause(me)
        if me == 0:
            counter, hub_round = int(hub[PAD_FGM + 0]), int(hub[PAD_FGM + 3])
            if (drain and hub_round == rnd and seen_round == rnd
                    and all(int(hub[PAD_VOTE + i]) == drain for i in range(k))):
                for p in range(k):
                    w.pads[p][PAD_FGM + 5] = drain     # everybody is done and no round is open: terminate
            elif counter > k and hub_round == rnd and seen_round == rnd:
                psi = float(sum(hub[PAD_PHI + i] for i in range(k)))
                if psi >= eps_psi * k * phi0:
                    for p in range(k):
                        w.pads[p][PAD_FGM + 4] = self.step
                    for p in range(k):
                        w.pads[p][PAD_FGM + 3] = rnd + 1      # round end, visible to every rank
                else:
                    hub[PAD_FGM + 2] = -psi / (2.0 * k)
                    hub[PAD_FGM + 0] = 0
                    hub[PAD_FGM + 1] = sub + 1
        if not act:
            return False
        # ---- round end: synchronous all-reduce keyed by the round number
        seq_entry, seq_exit = 2 * rnd + 1, 2 * rnd + 2
        self._signal_all(PAD_ENTRY, seq_entry)
        self._wait_all(PAD_ENTRY, seq_entry)
        lo, hi = w.shard(me)
        acc = scale * sum(w.bufs[p][lo:hi] for p in range(k))
        for p in range(k):
            w.bufs[p][lo:hi] = acc
        if me == 0:
            hub[PAD_FGM + 0] = 0
            hub[PAD_FGM + 2] = -phi0 / 2.0
            hub[PAD_PHI:PAD_PHI + k] = phi0
            hub[PAD_FGM + 1] = int(hub[PAD_FGM + 1]) + 1
        self._signal_all(PAD_EXIT, seq_exit)
        self._wait_all(PAD_EXIT, seq_exit)
        if mode == 0:
            self.e[:] = self.buf
        else:
            self.e += self.buf
            self.buf[:] = 0.0
        st["round"], st["sub"], st["sent"] = rnd + 1, -1, 0
        st["rounds"] += 1
        return True

    def fgm_flush(self, threshold: float, gen: int, eps_psi: float = 0.01, mode: int = 0, max_steps: int = 100000) -> int:
        """End of stream (FGMDriver.flush): keep stepping in drain mode - joining any round that is still
        announced - until the hub's terminate word of this generation arrives.  Returns the steps used."""
        for i in range(max_steps):
            self.fgm_step(threshold, eps_psi, mode, drain=gen)
            if int(self.w.pads[self.me][PAD_FGM + 5]) == gen:
                return i + 1
            time.sleep(0)
        self.w.pads[self.me][PAD_STATUS] = 4
        return max_steps

    def fgm_pending(self) -> bool:
        return int(self.w.pads[self.me][PAD_FGM + 3]) != self.fgm["round"]

    # ---- design model: reduce-scatter fused into the producer (ROADMAP item 0) -----------------------
    def fused_scatter_step(self, delta: np.ndarray) -> None:
        """One ORR-like step of the planned GEMM+reduce-scatter fusion, as a protocol model:
        (1) the producer (the GEMM epilogue) adds this rank's delta, shard by shard, straight into the shard
            OWNER's rotating buffer t mod 3 (remote reduce-add, no local copy);
        (2) ONE flag barrier;
        (3) every rank adds each owner's reduced shard to its private global vector `e`;
        (4) the owner clears buffer (t+2) mod 3 - the one read in step t-1 - which is safe exactly because the
            barrier of step t proves that every peer has finished step t-1, and nobody accumulates into that
            buffer before the barrier of step t+1, which needs this rank's arrival.
        `w.rot[r]` (3 x n per rank, shard-owner rows only are used) is created on first use."""
        w, me, k = self.w, self.me, self.w.k
        with w.lock:
            if not hasattr(w, "rot"):
                w.rot = [np.zeros((3, w.n)) for _ in range(k)]
        self.epoch += 1
        t = self.epoch
        cur, nxt = t % 3, (t + 2) % 3
        for owner in range(k):                         # (1) tile-by-tile remote reduce-add
            lo, hi = w.shard(owner)
            w.pause(me)
            with w.lock:
                w.rot[owner][cur, lo:hi] += delta[lo:hi]
        self._signal_all(PAD_ENTRY, t)                 # (2) the only barrier of the step
        self._wait_all(PAD_ENTRY, t)
        for owner in range(k):                         # (3) gather-add of the reduced shards
            lo, hi = w.shard(owner)
            w.pause(me)
            self.e[lo:hi] += w.rot[owner][cur, lo:hi]
        lo, hi = w.shard(me)                           # (4) clear the buffer of step t-1 for step t+2
        w.pause(me)
        w.rot[me][nxt, lo:hi] = 0.0

    # ---- barrier / query ---------------------------------------------------------------------------
    def barrier(self) -> None:
        """Own flag words and own counter (sync.cu: L_QEPOCH, QENTRY / QEXIT): independent of all-reduce epochs and
        FGM rounds."""
        self.qepoch += 1
        self._signal_all(PAD_QENTRY, 2 * self.qepoch + 1)
        self._wait_all(PAD_QENTRY, 2 * self.qepoch + 1)
        self._signal_all(PAD_QEXIT, 2 * self.qepoch + 2)
        self._wait_all(PAD_QEXIT, 2 * self.qepoch + 2)

    def query_reduce(self, vals) -> np.ndarray:
        self.qepoch += 1
        v = np.zeros(8)
        v[:len(vals)] = vals
        for p in range(self.w.k):
            self.w.pads[p][PAD_QUERY + 8 * self.me:PAD_QUERY + 8 * self.me + 8] = v
        self._signal_all(PAD_QENTRY, 2 * self.qepoch + 1)
        self._wait_all(PAD_QENTRY, 2 * self.qepoch + 1)
        out = sum(self.w.pads[self.me][PAD_QUERY + 8 * p:PAD_QUERY + 8 * p + 8].copy() for p in range(self.w.k))
        self._signal_all(PAD_QEXIT, 2 * self.qepoch + 2)
        self._wait_all(PAD_QEXIT, 2 * self.qepoch + 2)
        return out[:len(vals)]
