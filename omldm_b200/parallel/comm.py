"""Symmetric-memory arena for the fused protocol kernels.

One process per GPU; every rank allocates the same sequence of regions, so a region has the same
layout on every GPU and a peer's copy of any sub-buffer is `peer_base + offset`.  The peer
mappings come from CUDA IPC over a dedicated cudaMalloc (P2P over NVLink / NVSwitch); optionally
(`OMLDM_USE_SYMM_MEM=1`) from torch's symmetric memory, which also yields an NVLS multicast
alias used by the multimem path of csrc/sync.cu.  `torch.distributed` (NCCL or gloo) is used only
to exchange the handles.

This replaces the reference's transports: the Flink keyBy shuffle for worker->hub messages and
the Kafka `psMessages` feedback topic for hub->worker messages (network/FlinkNetwork.scala:242-293,
Job.scala:77-87,136-142).

Three modes:
  real     - world > 1, one rank per process / GPU (IPC or symmetric memory)
  single   - world == 1
  emulate  - one process plays k ranks on ONE GPU with a single cooperative launch per step; used
             by the single-GPU tests of the multi-rank kernels (B200_PROFILING.md: never run
             kernels that wait on each other as separate launches on one GPU)
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

import torch

from ..ops import _lib
from ..ops.functional import CommStruct

PAD_WORDS = 512
LOCAL_WORDS = 64
# pad word offsets (must match include/omldm_kernels.h)
PAD_ENTRY, PAD_EXIT, PAD_CLOCK, PAD_VOTE, PAD_PHI, PAD_FGM, PAD_STATUS, PAD_QUERY = 0, 16, 32, 48, 64, 80, 96, 128
PAD_HEART = 100
PAD_QENTRY, PAD_QEXIT, PAD_GSTEP, PAD_GENTRY, PAD_GEXIT = 256, 272, 288, 304, 320


class _RawCuda:
    """Expose a raw device allocation to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 3, "strides": None}


def _align(n: int, a: int = 1024) -> int:
    return (n + a - 1) // a * a


class Region:
    """One symmetric allocation: local bytes + the base pointer of every rank's copy."""

    def __init__(self, nbytes: int, device: torch.device, rank: int, world: int, group=None, emulate: int = 1):
        self.nbytes, self.device, self.rank, self.world, self.emulate = nbytes, device, rank, world, emulate
        self.mc_base = 0
        self._owned: List[int] = []
        self._imported: List[int] = []
        self._keep = []
        k = _lib.kernels()
        if device.type != "cuda":
            # CPU stand-in (host-side tests): plain tensors, "peers" are the emulated copies
            n = max(emulate, 1)
            self.locals = [torch.zeros(nbytes, dtype=torch.uint8) for _ in range(n)]
            self.bases = [t.data_ptr() for t in self.locals]
            return
        if emulate <= 1 and world > 1 and os.environ.get("OMLDM_USE_SYMM_MEM", "0") == "1":
            if self._try_symm_mem(nbytes, device, group):
                return
        n_local = emulate if emulate > 1 else 1
        self.locals = []
        bases_local = []
        for _ in range(n_local):
            out = ctypes.c_void_p()
            _lib.check(k.omldm_dev_malloc(ctypes.c_size_t(nbytes), ctypes.byref(out)), "arena malloc")
            self._owned.append(out.value)
            self.locals.append(torch.as_tensor(_RawCuda(out.value, nbytes), device=device))
            bases_local.append(out.value)
        if emulate > 1 or world == 1:
            self.bases = bases_local
            return
        import torch.distributed as dist
        handle = (ctypes.c_ubyte * 64)()
        _lib.check(k.omldm_ipc_export(ctypes.c_void_p(bases_local[0]), handle), "ipc export")
        gathered: List[Optional[bytes]] = [None] * world
        dist.all_gather_object(gathered, bytes(handle), group=group)
        self.bases = []
        for r in range(world):
            if r == rank:
                self.bases.append(bases_local[0])
                continue
            out = ctypes.c_void_p()
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(gathered[r])
            _lib.check(k.omldm_ipc_import(buf, ctypes.byref(out)), f"ipc import from rank {r}")
            self._imported.append(out.value)
            self.bases.append(out.value)
        dist.barrier(group=group)

    def _try_symm_mem(self, nbytes: int, device, group) -> bool:
        """torch symmetric memory: peer pointers plus an NVLS multicast alias (multimem path)."""
        try:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm_mem
            g = group if group is not None else dist.group.WORLD
            t = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
            t.zero_()
            hdl = symm_mem.rendezvous(t, g.group_name if hasattr(g, "group_name") else g)
            self.locals = [t]
            self.bases = [int(p) for p in hdl.buffer_ptrs]
            self.mc_base = int(getattr(hdl, "multicast_ptr", 0) or 0)
            self._keep = [hdl]
            torch.cuda.synchronize(device)
            dist.barrier(group=group)
            return True
        except Exception as e:   # fall back to CUDA IPC
            import warnings
            warnings.warn(f"symmetric memory unavailable ({e}); using CUDA IPC")
            return False

    def close(self) -> None:
        if self.device.type != "cuda":
            return
        k = _lib.kernels()
        for p in self._imported:
            k.omldm_ipc_close(ctypes.c_void_p(p))
        self.locals = []
        for p in self._owned:
            k.omldm_dev_free(ctypes.c_void_p(p))
        self._imported, self._owned = [], []


class PipelineComm:
    """pad | buf[n] | hub[n] for one pipeline, plus the pointer tables the kernels take."""

    def __init__(self, n_floats: int, device, rank: int = 0, world: int = 1, group=None, emulate: int = 1,
                 with_hub: bool = True, hub_floats: Optional[int] = None, n_hubs: int = 0):
        self.device = torch.device(device)
        self.rank, self.world, self.emulate = rank, world, emulate
        self.k = emulate if emulate > 1 else world
        self.n = (n_floats + 3) // 4 * 4
        pad_bytes = _align(PAD_WORDS * 4)
        buf_bytes = _align(self.n * 4)
        self.with_hub = with_hub
        self.hub_n = self.n if hub_floats is None else (hub_floats + 3) // 4 * 4     # the hub region may be larger
        hub_bytes = _align(self.hub_n * 4)
        nbytes = pad_bytes + buf_bytes + (hub_bytes if with_hub else 0)
        self.region = Region(nbytes, self.device, rank, world, group, emulate)
        self._off_buf, self._off_hub = pad_bytes, pad_bytes + buf_bytes
        nl = len(self.region.locals)
        self.pads = [self.region.locals[i][:PAD_WORDS * 4].view(torch.int32) for i in range(nl)]
        self.bufs = [self.region.locals[i][self._off_buf:self._off_buf + self.n * 4].view(torch.float32)
                     for i in range(nl)]
        self.hubs = [self.region.locals[i][self._off_hub:self._off_hub + self.hub_n * 4].view(torch.float32)
                     for i in range(nl)] if with_hub else [None] * nl
        # private (non-symmetric) words
        self.local = torch.zeros(LOCAL_WORDS * max(self.k, 1), dtype=torch.int32, device=self.device)
        self.epoch = 0
        self.clock = 0
        self.step = 0
        bases = self.region.bases
        ptr_pads = [b for b in bases]
        ptr_bufs = [b + self._off_buf for b in bases]
        ptr_hubs = [b + (self._off_hub if with_hub else self._off_buf) for b in bases]
        self.tables = torch.tensor([ptr_bufs, ptr_hubs, ptr_pads], dtype=torch.int64).to(self.device)
        self.struct = CommStruct(
            self.tables[0].data_ptr(), self.tables[1].data_ptr(), self.tables[2].data_ptr(),
            self.local.data_ptr(), (self.region.mc_base + self._off_buf) if self.region.mc_base else 0,
            rank if emulate <= 1 else 0, self.k, emulate if emulate > 1 else 0,
            n_hubs if 0 < n_hubs < self.k else 0)
        self.n_hubs = n_hubs if 0 < n_hubs < self.k else self.k
        self.multicast = bool(self.region.mc_base)

    # local views of this rank (rank 0 in emulate mode)
    @property
    def buf(self) -> torch.Tensor:
        return self.bufs[0]

    @property
    def hub(self) -> Optional[torch.Tensor]:
        return self.hubs[0]

    @property
    def pad(self) -> torch.Tensor:
        return self.pads[0]

    def struct_ptr(self):
        return ctypes.byref(self.struct)

    def next_epoch(self) -> int:
        self.epoch += 1
        return self.epoch

    def status(self) -> int:
        """Timeout bits recorded by the kernels (0 = healthy)."""
        return int(self.pad[PAD_STATUS].item())

    def heartbeat(self, beat: int, stream_ptr=None) -> None:
        """Publish this rank's liveness counter in every peer's pad (csrc/sync.cu::heartbeat_kernel; never waits)."""
        if self.k <= 1 or not (self.device.type == "cuda" or _lib._emulated):
            return
        _lib.check(_lib.kernels().omldm_heartbeat(self.struct_ptr(), ctypes.c_uint32(beat & 0xFFFFFFFF),
                                                  stream_ptr if stream_ptr is not None else _lib.stream_ptr()),
                   "heartbeat")

    def heartbeats(self) -> List[int]:
        """Last counter seen from every rank (own pad; index = rank)."""
        return [int(v) & 0xFFFFFFFF for v in self.pad[PAD_HEART:PAD_HEART + self.k].tolist()]

    def shard_bounds(self, r: int):
        """Float range of hub shard r (empty for ranks that own none, r >= n_hubs)."""
        n4 = self.n // 4
        per = (n4 + self.n_hubs - 1) // self.n_hubs
        return min(self.n, 4 * per * r), min(self.n, 4 * per * (r + 1))

    def close(self) -> None:
        self.pads, self.bufs, self.hubs = [], [], []
        self.region.close()
