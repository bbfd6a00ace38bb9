"""Device-side protocol drivers: one fused kernel launch per synchronisation step.

Each driver is the GPU counterpart of a worker/PS class pair of `host_protocols.py`
(utils/generators/MLNodeGenerator.scala:25-36,54-65).  `impl="fused"` launches the hand-written
kernels of csrc/sync.cu on the symmetric arena; `impl="nccl"` is the torch.distributed baseline
(all_reduce etc.) the fused path is measured against - "a path that only calls NCCL is the
baseline, not the product".

Buffer conventions (see csrc/sync.cu):
  mode 0 (merge "mean"): comm.buf IS the model w; e = last global estimate.
  mode 1 (merge "sum") : comm.buf is the additive delta since the last sync; e = global statistics.
"""
from __future__ import annotations

import ctypes
import math
from typing import Optional

import torch

from ..ops import _lib
from .comm import PipelineComm
from .host_protocols import cfg, cfg_float, cfg_int

_i, _f, _u = ctypes.c_int, ctypes.c_float, ctypes.c_uint32
_sz = ctypes.c_size_t


class ProtocolCounters:
    """models / bytes / blocks shipped, as in FlinkHub.scala:118-127."""

    def __init__(self, protocol: str, max_msg_params: int = 2000):
        self.protocol = protocol
        self.models = self.bytes = self.blocks = self.syncs = 0
        self.max_msg_params = max_msg_params

    def ship(self, n_params: int, k: int) -> None:
        self.syncs += 1
        self.models += 2 * k                      # k pushes + k installs
        self.bytes += 2 * k * 4 * n_params
        self.blocks += 2 * k * max(1, math.ceil(n_params / self.max_msg_params))


class SyncDriver:
    protocol = "CentralizedTraining"

    @property
    def graph_safe(self) -> bool:
        """True if `step()` enqueues device work only (no host reads), so it can be graph-captured."""
        return self.impl == "fused" or self.protocol in ("CentralizedTraining", "SingleLearner", "Synchronous",
                                                         "Asynchronous", "SSP", "EASGD")

    def __init__(self, comm: PipelineComm, e: Optional[torch.Tensor], merge_mode: str, tc: Optional[dict],
                 impl: str = "fused", group=None, max_msg_params: int = 2000):
        self.comm, self.e, self.tc = comm, e, tc or {}
        self.mode = 1 if merge_mode == "sum" else 0
        self.k = comm.k
        self.scale = 1.0 if self.mode == 1 else 1.0 / self.k
        # host tensors have no fused path - except under the CPU emulation of the kernels with k emulated ranks
        # (tools/cuda_emu), which is how these driver classes are tested without GPUs
        self.impl = impl if (comm.device.type == "cuda" or (_lib._emulated and comm.emulate > 1)) else "nccl"
        self.group = group
        self.sync_every = cfg_int(tc, ["syncEvery", "syncPeriod"], 1)
        self.blocks = cfg_int(tc, ["syncBlocks"], 0)
        self.counters = ProtocolCounters(self.protocol, max_msg_params)
        self._batches = 0
        self.synced_last_step = False

    # ------------------------------------------------------------------ helpers
    def _dist(self):
        import torch.distributed as dist
        return dist if (dist.is_available() and dist.is_initialized() and self.comm.world > 1) else None

    def _eptr(self):
        return _lib.ptr(self.e)

    def after_batch(self) -> bool:
        """Called after every learner mini-batch; returns True if the model may have changed."""
        due = self.host_tick()
        self.synced_last_step = bool(self.step()) if due else False
        return self.synced_last_step

    def host_tick(self) -> bool:
        """Host bookkeeping of one mini-batch; True if a protocol step is due now."""
        self._batches += 1
        if self._batches % self.sync_every:
            return False
        self.account()
        return True

    def account(self) -> None:
        """models / bytes / blocks shipped by one protocol step (FlinkHub.scala:118-127)."""

    def step(self) -> bool:
        return False

    def flush(self) -> None:
        """End of stream: make sure no rank is left waiting inside a collective."""

    def _nccl_allreduce(self) -> None:
        d = self._dist()
        buf = self.comm.buf
        if d is not None:
            d.all_reduce(buf, group=self.group)
        if self.mode == 0:
            if self.scale != 1.0:
                buf.mul_(self.scale)
            if self.e is not None:
                self.e.copy_(buf)
        else:
            self.e.add_(buf)
            buf.zero_()

    def _fused_allreduce(self) -> None:
        c = self.comm
        _lib.check(_lib.kernels().omldm_sync_allreduce(c.struct_ptr(), self._eptr(), _sz(c.n), _f(self.scale),
                                                       _i(self.mode), _u(0), _i(self.blocks),
                                                       _lib.stream_ptr()), "sync_allreduce")


class LocalDriver(SyncDriver):
    """CentralizedTraining: a lone learner (parallelism 1, FlinkSpoke.scala:213-216)."""

    def step(self) -> bool:
        # nothing to reconcile: additive learners always read (e + buf), so the delta may simply keep
        # accumulating in buf - no fold pass per mini-batch
        return False


class SynchronousDriver(SyncDriver):
    protocol = "Synchronous"

    def account(self):
        self.counters.ship(self.comm.n, self.k)

    def step(self) -> bool:
        self._fused_allreduce() if self.impl == "fused" else self._nccl_allreduce()
        return True


class PipelinedSynchronousDriver(SyncDriver):
    """Synchronous protocol for ADDITIVE statistics with the all-reduce hidden behind the next
    mini-batch.  The learner accumulates step t into delta buffer t % 2 (the two symmetric buffers
    of the region: `buf` and `hub`).  At the start of step t+1 the fused all-reduce of buffer t % 2
    is forked onto a communication stream and runs beside the holdout / Gram kernels of step t+1,
    which write the other buffer; the streams join at the end of the step, so the whole step is
    still one capturable unit (one CUDA graph per batch buffer and parity).  Every contribution is
    folded into the global statistics exactly once - the sum is order independent - it just becomes
    visible to the peers one step later; `flush()` pushes the last delta at end of stream.  Not valid
    for averaged models."""
    protocol = "Synchronous"

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        if self.mode != 1:
            raise ValueError("pipelined synchronisation needs additive statistics (merge mode 'sum')")
        c = self.comm
        self.stream = torch.cuda.Stream(device=c.device)
        self.cur = 0                   # delta buffer the learner writes in the current step
        self.bufs = [c.buf, c.hub]
        from ..ops.functional import CommStruct
        # second view of the region: the kernels' "buf" table points at the hub buffers
        self._struct_b = CommStruct(c.tables[1].data_ptr(), c.tables[0].data_ptr(), c.tables[2].data_ptr(),
                                    c.local.data_ptr(), 0, c.struct.rank, c.struct.world, c.struct.emulate, c.struct.nhub)
        self.on_switch = None          # learner callback: (index) -> use that delta buffer from now on
        if self.blocks == 0:
            self.blocks = 48           # leave the SMs to the tensor-core kernel it overlaps with
        self._forked = False

    def graph_key(self):
        return self.cur

    def _launch_sync(self, i: int) -> None:
        c = self.comm
        struct = ctypes.byref(c.struct) if i == 0 else ctypes.byref(self._struct_b)
        _lib.check(_lib.kernels().omldm_sync_allreduce(struct, self._eptr(), _sz(c.n), _f(self.scale), _i(1),
                                                       _u(0), _i(self.blocks), _lib.stream_ptr()), "sync_allreduce")

    def begin_step(self) -> None:
        """Fork: all-reduce the delta written in the previous step while this step computes."""
        main = torch.cuda.current_stream(self.comm.device)
        self.stream.wait_stream(main)
        with torch.cuda.stream(self.stream):
            self._launch_sync(self.cur ^ 1)
        self._forked = True

    def step(self) -> bool:
        """Join the communication stream (end of the step's device work)."""
        if self._forked:
            torch.cuda.current_stream(self.comm.device).wait_stream(self.stream)
            self._forked = False
        return False                   # derived state is refreshed by the learner's own cadence

    def host_tick(self) -> bool:
        due = super().host_tick()
        self.cur ^= 1
        if self.on_switch is not None:
            self.on_switch(self.cur)
        return due

    def drain(self) -> None:
        """Nothing is in flight between steps (every step joins); kept for the learner's snapshot."""

    def flush(self) -> None:
        """End of stream: the delta of the last step has not been shared yet."""
        self._launch_sync(self.cur ^ 1)

    def account(self):
        self.counters.ship(self.comm.n, self.k)


class ScatterSynchronousDriver(SyncDriver):
    """Synchronous protocol for ADDITIVE statistics with the reduce-scatter moved into the producer (opt-in
    `syncFusion: "scatter"`; csrc/sync.cu::sync_scatter_push_kernel / sync_scatter_gather_kernel, protocol modelled
    in parallel/fake_backend.py::fused_scatter_step).  The learner accumulates into a PRIVATE delta; the symmetric
    region holds three rotating copies into which every rank's delta is reduced at the shard owners without a
    barrier, then one flag barrier and a gather-add into the private global vector - one barrier per step instead
    of two and no separate reduce pass.  The push is a stand-alone kernel today; it disappears once the Gram GEMM's
    TMA reduce-add epilogue addresses the owners' buffers itself (ROADMAP item 0)."""
    protocol = "Synchronous"

    def __init__(self, comm, e, merge_mode, tc, impl="fused", group=None, max_msg_params=2000, delta=None):
        super().__init__(comm, e, merge_mode, tc, impl, group, max_msg_params)
        if self.mode != 1 or delta is None or e is None:
            raise ValueError("the scatter-fused synchronisation needs additive statistics, a private delta and e")
        if comm.n % 3 or (comm.n // 3) % 4 or delta.numel() != (comm.n // 3) * max(comm.emulate, 1):
            raise ValueError("the symmetric region must hold three copies of the delta")
        self.n = comm.n // 3
        self.delta = delta
        # `syncFusion: "gemm"`: the learner's GEMM reduces into the owners' copies itself (gemm_outer_scatter), the push
        # kernel disappears.  The rotating copy is then chosen on the HOST (it is baked into the GEMM's tensor maps):
        # `hstep` counts protocol steps, one CUDA graph per hstep mod 3 (`graph_key`).
        self.gemm_fused = (str(cfg(tc, ["syncFusion"], "none")).lower() == "gemm" and self.impl == "fused")
        self.hstep = 1

    def account(self):
        self.counters.ship(self.n, self.k)

    def graph_key(self):
        return self.hstep % 3 if self.gemm_fused else 0

    def peer_copy_ptrs(self):
        """Base addresses (one per rank) of the rotating copy that the current step accumulates into."""
        off = 4 * self.n * (self.hstep % 3)
        return [int(b) + self.comm._off_buf + off for b in self.comm.region.bases]

    def host_tick(self) -> bool:
        due = super().host_tick()
        if due:
            self.hstep += 1
        return due

    def step(self) -> bool:
        c = self.comm
        if self.impl == "fused":
            k = _lib.kernels()
            if self.gemm_fused:
                # explicit rotation step (non-zero, congruent to hstep mod 3) so that a captured launch stays valid
                # for every later step with the same hstep mod 3; the epoch stays device managed
                _lib.check(k.omldm_sync_scatter_gather(c.struct_ptr(), self._eptr(), _sz(self.n), _f(self.scale), _u(0),
                                                       _u(3 + self.hstep % 3), _i(self.blocks), _lib.stream_ptr()),
                           "sync_scatter_gather")
                return True
            _lib.check(k.omldm_sync_scatter_push(c.struct_ptr(), _lib.ptr(self.delta), _sz(self.n), _u(0),
                                                 _i(self.blocks), _lib.stream_ptr()), "sync_scatter_push")
            _lib.check(k.omldm_sync_scatter_gather(c.struct_ptr(), self._eptr(), _sz(self.n), _f(self.scale), _u(0),
                                                   _u(0), _i(self.blocks), _lib.stream_ptr()), "sync_scatter_gather")
            return True
        d = self._dist()
        if d is not None:
            d.all_reduce(self.delta, group=self.group)
        self.e.add_(self.delta, alpha=self.scale)
        self.delta.zero_()
        return True


class ShardedGramDriver(SyncDriver):
    """Synchronous protocol for the ORR Gram statistics with the hub SHARDED over the GPUs and the worker -> hub push
    performed by the kernel that computes the update (csrc/gram_f16.cu; csrc/sync.cu::sync_gram_push_kernel for the
    tf32 path).  Every worker keeps its own running sum `acc` (private) and, each step, copies its rows into
    ITS slot at the GPU that owns them (`hub` region = k slot matrices; single writer per slot, so plain stores over
    NVLink: no remote atomics, no flags, nothing waits).  `hubPush: "lead"` (default): the Gram kernel ships the
    running sum as it was before the step while its MMAs run - the transfer costs nothing, the slots trail by one
    mini-batch and `gather()` first brings them up to date; `hubPush: "tail"`: the last CTA of every half tile ships the
    step's result, the slots are exact after every step (measured + 6 us per step at 2 GPUs).  Workers need the global statistics only when they refresh
    their weights (`solveEvery`) or answer a query, so the hub -> worker broadcast of FlinkNetwork.scala:282-291
    becomes a pull: `gather()` - entry rendezvous, every owner sums its k slots and stores the result into every
    worker's global matrix (`buf` region), exit rendezvous.  Default for Synchronous ORR on the fused path
    (`syncFusion: "hub"`)."""
    protocol = "Synchronous"

    @staticmethod
    def region_floats(ld: int, k: int):
        """(buf floats, hub floats) of the symmetric region: the global matrix and k slot matrices."""
        return ld * ld, max(k, 1) * ld * ld

    def __init__(self, comm, e, merge_mode, tc, impl="fused", group=None, max_msg_params=2000, ld: int = 0, n: int = 0):
        super().__init__(comm, e, merge_mode, tc, impl, group, max_msg_params)
        k = max(comm.k, 1)
        if self.mode != 1 or ld * ld != comm.n or not comm.with_hub or comm.hub_n < k * ld * ld:
            raise ValueError("the sharded hub needs additive statistics, an ld x ld global matrix and k slot matrices")
        from ..ops.functional import GramPeers, gram_f16_scratch
        self.ld, self.n_rows = ld, n
        h = comm.n_hubs                      # HubParallelism: only ranks 0 .. h-1 own rows
        rpo = (ld // h) // 32 * 32
        while rpo * h < n:
            rpo += 32
        self.rows_per_owner = rpo
        self.scratch = gram_f16_scratch(comm.device)
        # this worker's running sum; under emulation one private matrix per emulated rank
        self.acc = torch.zeros(max(comm.emulate, 1) * ld * ld, device=comm.device)
        self.peers = GramPeers()
        bases = comm.region.bases
        for p_ in range(k if comm.emulate <= 1 else 0):
            self.peers.owner_base[p_] = int(bases[p_]) + comm._off_hub + 4 * comm.rank * ld * ld
        self.peers.world, self.peers.rank, self.peers.rows_per_owner = k, comm.rank, rpo
        self.lead = str(cfg(tc, ["hubPush"], "lead")).lower() != "tail"
        self.peers.lead = 1 if self.lead else 0
        self.gathers = 0

    def account(self):
        self.counters.ship(self.n_rows * (self.n_rows + 1) // 2, self.k)

    def step(self) -> bool:
        return False                   # the push happened in the producer; derived state follows the learner's cadence

    def push(self) -> None:
        """Copy the running sum's upper blocks into this worker's slots (for running sums kept by the tf32 path)."""
        c = self.comm
        _lib.check(_lib.kernels().omldm_sync_gram_push(c.struct_ptr(), _lib.ptr(self.acc), _i(self.ld), _i(self.n_rows),
                                                       _i(self.rows_per_owner), _i(self.blocks), _lib.stream_ptr()),
                   "sync_gram_push")

    def gather(self) -> None:
        """Collective: after it, `comm.buf` holds the global statistics (the sum of every worker's running sum)."""
        c = self.comm
        self.gathers += 1
        if self.lead and c.emulate <= 1:
            self.push()                # the slots trail the running sum by one mini-batch: bring them up to date
        _lib.check(_lib.kernels().omldm_sync_gram_gather(c.struct_ptr(), _i(self.ld), _i(self.n_rows),
                                                         _i(self.rows_per_owner), _u(0), _i(self.blocks),
                                                         _lib.stream_ptr()), "sync_gram_gather")


def hub_mode(tc) -> str:
    return str(cfg(tc, ["syncFusion"], "hub")).lower()


def wants_sharded_hub(protocol: str, merge_mode: str, tc, world: int, impl: str, device_type: str) -> bool:
    """`syncFusion: "hub"` (the default for additive learners that support it): Synchronous on more than one GPU."""
    return (protocol == "Synchronous" and merge_mode == "sum" and world > 1 and impl == "fused" and device_type == "cuda"
            and hub_mode(tc) == "hub"
            and cfg_int(tc, ["syncEvery", "syncPeriod"], 1) == 1)


def wants_scatter(protocol: str, merge_mode: str, tc, world: int, impl: str) -> bool:
    """`syncFusion: "scatter"` / `"gemm"` apply to Synchronous pipelines of additive learners on more than one GPU."""
    return (protocol == "Synchronous" and merge_mode == "sum" and world > 1
            and str(cfg(tc, ["syncFusion"], "none")).lower() in ("scatter", "gemm")
            and cfg_int(tc, ["syncEvery", "syncPeriod"], 1) == 1)


class AsynchronousDriver(SyncDriver):
    protocol = "Asynchronous"

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        if self.e is None:
            raise ValueError("asynchronous protocols need the estimate vector e")
        # hub shards start at the initial global state (identical on every rank)
        if self.comm.with_hub and self.comm.hub is not None:
            for h in self.comm.hubs:
                h.copy_(self.e if self.mode == 1 else self.comm.buf)

    def _fused(self, ssp: bool = False, staleness: int = 0) -> None:
        c = self.comm
        k = _lib.kernels()
        if ssp:
            _lib.check(k.omldm_sync_ssp(c.struct_ptr(), self._eptr(), _sz(c.n), _f(self.scale), _i(self.mode),
                                        _u(0), _i(staleness), _i(self.blocks), _lib.stream_ptr()), "sync_ssp")
        else:
            _lib.check(k.omldm_sync_async(c.struct_ptr(), self._eptr(), _sz(c.n), _f(self.scale), _i(self.mode),
                                          _i(self.blocks), _lib.stream_ptr()), "sync_async")

    def account(self):
        self.counters.ship(self.comm.n, 1)

    def step(self) -> bool:
        if self.impl == "fused":
            self._fused()
        else:   # closest NCCL equivalent: everybody folds everybody's delta in lock-step
            self._nccl_delta_allreduce()
        return True

    def _nccl_delta_allreduce(self) -> None:
        d = self._dist()
        buf = self.comm.buf
        if self.mode == 0:
            delta = (buf - self.e) * self.scale
            if d is not None:
                d.all_reduce(delta, group=self.group)
            self.e.add_(delta)
            buf.copy_(self.e)
        else:
            if d is not None:
                d.all_reduce(buf, group=self.group)
            self.e.add_(buf)
            buf.zero_()


class SSPDriver(AsynchronousDriver):
    protocol = "SSP"

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.staleness = cfg_int(self.tc, ["staleness", "s", "stalenessBound"], 3)

    def step(self) -> bool:
        if self.impl == "fused":
            self._fused(ssp=True, staleness=self.staleness)
        else:
            self._nccl_delta_allreduce()
        return True


class EASGDDriver(SyncDriver):
    protocol = "EASGD"

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.alpha = cfg_float(self.tc, ["alpha", "elasticity", "rho"], 0.1)
        self.sync_every = cfg_int(self.tc, ["tau", "commPeriod", "syncEvery"], 1)
        if self.mode == 1:
            raise ValueError("EASGD needs an averaged model (merge mode 'mean')")
        for h in self.comm.hubs:
            if h is not None:
                h.copy_(self.comm.buf)

    def step(self) -> bool:
        c = self.comm
        if self.impl == "fused":
            _lib.check(_lib.kernels().omldm_sync_easgd(c.struct_ptr(), _sz(c.n), _f(self.alpha), _i(self.blocks),
                                                       _lib.stream_ptr()), "sync_easgd")
        else:
            d = self._dist()
            # centre variable replicated in hub[]: x~ += alpha * sum_i (x_i - x~); x_i -= alpha (x_i - x~)
            diff = (c.buf - c.hub) * self.alpha
            c.buf.sub_(diff)
            if d is not None:
                d.all_reduce(diff, group=self.group)
            c.hub.add_(diff)
        return True

    def account(self):
        self.counters.ship(self.comm.n, 1)


class GMDriver(SyncDriver):
    protocol = "GM"

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.threshold = cfg_float(self.tc, ["threshold", "T", "gmThreshold"], 0.5)
        self.rounds = 0
        if self.e is None:
            raise ValueError("GM needs the estimate vector e")

    def step(self) -> bool:
        c = self.comm
        if self.impl == "fused":
            _lib.check(_lib.kernels().omldm_sync_gm(c.struct_ptr(), self._eptr(), _sz(c.n), _f(self.scale),
                                                    _i(self.mode), _f(self.threshold), _u(0),
                                                    _i(self.blocks), _lib.stream_ptr()), "sync_gm")
            return True   # whether a round happened is only known on the device
        d = self._dist()
        diff = c.buf - self.e if self.mode == 0 else c.buf
        flag = ((diff * diff).sum() > self.threshold).float().reshape(1)
        if d is not None:
            d.all_reduce(flag, op=d.ReduceOp.MAX, group=self.group)
        if float(flag.item()) > 0:
            self._nccl_allreduce()
            self.rounds += 1
            self.counters.ship(c.n, self.k)
            return True
        return False


class FGMDriver(SyncDriver):
    protocol = "FGM"

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.threshold = cfg_float(self.tc, ["threshold", "T", "fgmThreshold"], 0.5)
        self.eps_psi = cfg_float(self.tc, ["epsPsi", "eps_psi"], 0.01)
        if self.e is None:
            raise ValueError("FGM needs the estimate vector e")
        n_state = 8 * max(self.comm.emulate, 1)
        self.state = torch.zeros(n_state, dtype=torch.float32, device=self.comm.device)
        self.state[1::8] = -1.0   # force re-anchoring on the first step
        self.rounds = 0
        self.flush_mode = str(cfg(self.tc, ["fgmFlush"], "drain")).lower()
        self.flush_timeout_s = cfg_float(self.tc, ["fgmFlushTimeout"], 5.0)
        self.flush_timeouts = 0
        self._drain_gen = 0
        if self.impl == "fused":
            # hub (rank 0) starts the first round: theta = sqrt(T)/2, every phi = phi(0)
            phi0 = -math.sqrt(self.threshold)
            from .comm import PAD_FGM, PAD_PHI
            pad0 = self.comm.pads[0]
            fl = pad0.view(torch.float32)
            fl[PAD_FGM + 2] = -phi0 / 2.0
            fl[PAD_PHI:PAD_PHI + 16] = phi0

    def step(self) -> bool:
        c = self.comm
        if self.impl == "fused":
            _lib.check(_lib.kernels().omldm_sync_fgm(c.struct_ptr(), self._eptr(), _sz(c.n), _f(self.scale),
                                                     _i(self.mode), _f(self.threshold), _f(self.eps_psi), _u(0),
                                                     _lib.ptr(self.state), _i(self.blocks), _lib.stream_ptr()),
                       "sync_fgm")
            return True
        d = self._dist()
        diff = c.buf - self.e if self.mode == 0 else c.buf
        phi0 = -math.sqrt(self.threshold)
        psi = (torch.sqrt((diff * diff).sum()) + phi0).reshape(1)
        if d is not None:
            d.all_reduce(psi, group=self.group)
        if float(psi.item()) >= self.eps_psi * self.k * phi0:
            self._nccl_allreduce()
            self.rounds += 1
            self.counters.ship(c.n, self.k)
            return True
        return False

    def device_rounds(self) -> int:
        return int(self.state[4].item()) if self.impl == "fused" else self.rounds

    def flush(self) -> None:
        """End of stream (collective).  The round-end all-reduce waits for every rank inside the kernel, so a
        rank must not simply stop stepping while a peer may still announce a round.  Drain handshake
        (csrc/sync.cu, modelled and stress-tested in parallel/fake_backend.py): keep stepping in drain mode -
        which joins any round still announced - until the hub, having heard "no more data" from everybody with no
        round open, writes this flush generation into every pad.  Works for any skew and for unequal step counts."""
        if self.impl != "fused" or self.comm.k <= 1:
            return
        if self.flush_mode == "legacy":
            return self._flush_legacy()
        from .comm import PAD_FGM
        import time
        self._drain_gen += 1
        c, g = self.comm, self._drain_gen
        deadline = time.monotonic() + self.flush_timeout_s
        while True:
            _lib.check(_lib.kernels().omldm_sync_fgm_drain(c.struct_ptr(), self._eptr(), _sz(c.n), _f(self.scale),
                                                           _i(self.mode), _f(self.threshold), _f(self.eps_psi), _u(0),
                                                           _lib.ptr(self.state), _u(g), _i(self.blocks),
                                                           _lib.stream_ptr()), "sync_fgm_drain")
            if int(c.pad[PAD_FGM + 5].item()) == g:      # .item() waits for the step just launched
                return
            if time.monotonic() > deadline:
                import warnings
                warnings.warn("FGM flush: no terminate word from the hub within %.1f s" % self.flush_timeout_s)
                self.flush_timeouts += 1
                return
            time.sleep(2e-5)

    def _flush_legacy(self) -> None:
        """Round-1 flush: agree on the host whether a round is pending anywhere and, if so, step once more on
        every rank.  Needs equal step counts and small skew between ranks."""
        from .comm import PAD_FGM
        d = self._dist()
        for _ in range(4):
            torch.cuda.synchronize(self.comm.device)
            pending = int(self.comm.pad[PAD_FGM + 3].item()) != int(self.state[0].item())
            flag = torch.tensor([1.0 if pending else 0.0], device=self.comm.device)
            if d is not None:
                d.all_reduce(flag, op=d.ReduceOp.MAX, group=self.group)
            if float(flag.item()) == 0.0:
                return
            self.step()


DRIVERS = {"CentralizedTraining": LocalDriver, "SingleLearner": LocalDriver, "Synchronous": SynchronousDriver,
           "Asynchronous": AsynchronousDriver, "SSP": SSPDriver, "EASGD": EASGDDriver, "GM": GMDriver,
           "FGM": FGMDriver}


def make_driver(protocol: str, comm: PipelineComm, e, merge_mode: str, tc, impl: str = "fused", group=None,
                max_msg_params: int = 2000, pipelined_ok: bool = False, scatter_delta=None) -> SyncDriver:
    if scatter_delta is not None:
        drv = ScatterSynchronousDriver(comm, e, merge_mode, tc, impl, group, max_msg_params, delta=scatter_delta)
        drv.counters.protocol = protocol
        return drv
    cls = DRIVERS.get(protocol, AsynchronousDriver)
    if comm.k == 1:
        cls = LocalDriver
    overlap = str(cfg(tc, ["overlapSync"], "false")).lower() in ("1", "true", "yes")   # opt-in: measured slower at 2 GPUs
    if (cls is SynchronousDriver and merge_mode == "sum" and overlap and impl == "fused" and comm.k > 1
            and comm.device.type == "cuda" and comm.emulate <= 1 and cfg_int(tc, ["syncEvery", "syncPeriod"], 1) == 1
            and pipelined_ok):
        cls = PipelinedSynchronousDriver
    if cls is EASGDDriver and merge_mode == "sum":
        cls = AsynchronousDriver   # elastic averaging is meaningless for additive statistics
    drv = cls(comm, e, merge_mode, tc, impl, group, max_msg_params)
    drv.counters.protocol = protocol if comm.k > 1 else "CentralizedTraining"
    return drv


def device_barrier(comm: PipelineComm) -> None:
    """Termination probe rendezvous (StatisticsOperator.scala:135-142) on the signal pads."""
    if comm.device.type == "cuda" and comm.k > 1:
        _lib.check(_lib.kernels().omldm_device_barrier(comm.struct_ptr(), _u(0), _lib.stream_ptr()),
                   "device_barrier")


def split_count(c: int):
    """A record count as three 20-bit chunks: their sums over up to 16 ranks stay exact in fp32 (the query words are
    fp32; a count above 2^24 would not)."""
    c = int(c)
    return float(c & 0xFFFFF), float((c >> 20) & 0xFFFFF), float(c >> 40)


def join_count(lo: float, mid: float, hi: float) -> int:
    return int(round(lo)) + (int(round(mid)) << 20) + (int(round(hi)) << 40)


def query_reduce(comm: PipelineComm, vals: torch.Tensor, impl: str = "fused", group=None) -> torch.Tensor:
    """Sum of up to 8 query scalars over ranks (ResponseConstructor.scala:43-52); `vals` is padded to 8 words."""
    v8 = torch.zeros(8, dtype=torch.float32, device=vals.device)
    v8[: vals.numel()] = vals
    out = torch.empty_like(v8)
    if comm.device.type == "cuda" and comm.k > 1 and impl == "fused":
        _lib.check(_lib.kernels().omldm_query_reduce(comm.struct_ptr(), _lib.ptr(v8), _lib.ptr(out),
                                                     _u(0), _lib.stream_ptr()), "query_reduce")
        return out[: vals.numel()]
    out.copy_(v8)
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and comm.world > 1:
        if out.device.type == "cpu":
            o64 = out.double()          # gloo path: exact sums
            dist.all_reduce(o64, group=group)
            out = o64.float()
        else:
            dist.all_reduce(out, group=group)
    return out[: vals.numel()]
