"""Pinned host ring buffers ("topics") and synthetic stream producers (csrc/host/*.cpp).

A `Topic` replaces one Kafka topic of the reference (trainingData / forecastingData, Job.scala:42-67):
fixed-size slots of pinned host memory, each holding one mini-batch of staged record rows, filled by
native producer threads and drained by the worker, which copies a slot to its GPU with a single
async H2D transfer.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from ..ops import _lib
from ..ops.functional import f16s_pitch, pack_f16s, pack_wire_f16, pad32, wire_f16_pitch

KIND_BINARY, KIND_REGRESSION, KIND_CLUSTERS, KIND_MULTICLASS, KIND_SPARSE_MULTICLASS = 0, 1, 2, 3, 4
MODE_GENERATE, MODE_REPLAY = 0, 1

_configured = False


def _h():
    global _configured
    h = _lib.host()
    if not _configured:
        vp, i, i64, u64, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_uint64, ctypes.c_size_t
        h.omldm_ring_create.restype = vp
        h.omldm_ring_create.argtypes = [vp, i, sz]
        h.omldm_ring_destroy.argtypes = [vp]
        h.omldm_ring_acquire_write.argtypes = [vp, i]
        h.omldm_ring_commit_write.argtypes = [vp, i, i64]
        h.omldm_ring_acquire_read.argtypes = [vp, i, ctypes.POINTER(i64)]
        h.omldm_ring_release_read.argtypes = [vp, i]
        h.omldm_ring_close.argtypes = [vp]
        h.omldm_ring_produced.restype = i64
        h.omldm_ring_produced.argtypes = [vp]
        h.omldm_ring_consumed.restype = i64
        h.omldm_ring_consumed.argtypes = [vp]
        h.omldm_ring_slot_ptr.restype = vp
        h.omldm_ring_slot_ptr.argtypes = [vp, i]
        h.omldm_producer_start.restype = vp
        h.omldm_producer_start.argtypes = [vp, i, i, i, i, i, u64, i, i64, i]
        h.omldm_producer_start_wire.restype = vp
        h.omldm_producer_start_wire.argtypes = [vp, i, i, i, i, i, u64, i, i64, i, i]
        h.omldm_producer_start_from.restype = vp
        h.omldm_producer_start_from.argtypes = [vp, i, i, i, i, i, u64, i, i64, i, i, i64]
        h.omldm_producer_stop.argtypes = [vp]
        h.omldm_generate_batch_f16.argtypes = [vp, i, i, i, i, u64, u64]
        h.omldm_generate_batch_f16s.argtypes = [vp, i, i, i, i, u64, u64]
        h.omldm_generate_batch.argtypes = [vp, i, i, i, i, i, u64, u64]
        h.omldm_generate_sparse_batch.argtypes = [vp, i, i, i, i, u64, u64]
        h.omldm_parse_json_records.restype = i
        h.omldm_parse_json_records.argtypes = [ctypes.c_char_p, sz, i, vp, i, i, ctypes.POINTER(i),
                                               ctypes.POINTER(sz)]
        h.omldm_parse_json_records_mt.restype = i
        h.omldm_parse_json_records_mt.argtypes = [vp, sz, i, vp, i, i, ctypes.POINTER(i), ctypes.POINTER(sz), i]
        _configured = True
    return h


class Topic:
    """Ring of `n_slots` batches of `batch` staged rows with `d` features.

    `wire="f16"` keeps the slots in the compact wire format instead ([fp16 x d | pad | fp32 target],
    `wire_f16_pitch(d)` halves per record): half the bytes per record on the host->device link, expanded to
    staged fp32 rows on the GPU (`Engine.train_batch_host` recognises the dtype).
    `wire="f16s"` keeps fp16 STAGED rows ([x | 1 | y_hi | y_lo | 0-pad], `f16s_pitch(d)` halves): the layout the
    fp16 Gram kernel of ORR reads in place, so the slot is trained on as it lands on the GPU."""

    def __init__(self, d: int, batch: int, n_slots: int = 8, pin: Optional[bool] = None, wire: str = "f32"):
        if wire not in ("f32", "f16", "f16s"):
            raise ValueError("wire must be 'f32', 'f16' or 'f16s'")
        self.d, self.batch, self.n_slots, self.wire = d, batch, n_slots, wire
        self.ld = pad32(d + 2)
        if wire in ("f16", "f16s"):
            self.wire_ld = wire_f16_pitch(d) if wire == "f16" else f16s_pitch(d)
            self.storage = torch.zeros(n_slots, batch, self.wire_ld, dtype=torch.float16)
        else:
            self.wire_ld = self.ld
            self.storage = torch.zeros(n_slots, batch, self.ld, dtype=torch.float32)
        if pin is None:
            pin = torch.cuda.is_available()
        if pin:
            self.storage = self.storage.pin_memory()
        self.slot_bytes = batch * self.wire_ld * self.storage.element_size()
        self._ring = _h().omldm_ring_create(self.storage.data_ptr(), n_slots, self.slot_bytes)
        if not self._ring:
            raise RuntimeError("ring creation failed")
        self._producer = None

    # ---- producer side -------------------------------------------------------------------------
    def start_synthetic(self, kind: int, seed: int = 0, n_aux: int = 0, mode: int = MODE_GENERATE,
                        max_batches: int = -1, n_threads: int = 1, first_batch: int = 0) -> None:
        """`first_batch` > 0 resumes the stream at that position (batches first_batch .. max_batches - 1)."""
        self._producer = _h().omldm_producer_start_from(self._ring, kind, self.d, self.ld, self.batch, n_aux, seed,
                                                        mode, max_batches, n_threads,
                                                        {"f32": 0, "f16": 1, "f16s": 2}[self.wire], first_batch)
        if not self._producer:
            raise RuntimeError("producer start failed")

    def put(self, rows: torch.Tensor, timeout_ms: int = -1) -> bool:
        """Publish one batch of staged rows from Python (e.g. parsed JSON records)."""
        s = _h().omldm_ring_acquire_write(self._ring, timeout_ms)
        if s < 0:
            return False
        n = rows.shape[0]
        if self.wire == "f16" and rows.dtype != torch.float16:
            rows = pack_wire_f16(rows, self.d)
        elif self.wire == "f16s" and rows.dtype != torch.float16:
            rows = pack_f16s(rows, self.d)
        self.storage[s, :n].copy_(rows)
        if n < self.batch:
            self.storage[s, n:].zero_()
        _h().omldm_ring_commit_write(self._ring, s, n)
        return True

    def close(self) -> None:
        _h().omldm_ring_close(self._ring)

    # ---- consumer side -------------------------------------------------------------------------
    def get(self, timeout_ms: int = -1) -> Optional[Tuple[int, torch.Tensor, int]]:
        """(slot, pinned batch view, n_records) or None at end of stream / timeout."""
        n = ctypes.c_int64(0)
        while True:
            s = _h().omldm_ring_acquire_read(self._ring, timeout_ms, ctypes.byref(n))
            if s < 0:
                return None
            if n.value == 0:        # empty end-marker batch
                _h().omldm_ring_release_read(self._ring, s)
                continue
            return s, self.storage[s], int(n.value)

    def release(self, slot: int) -> None:
        _h().omldm_ring_release_read(self._ring, slot)

    @property
    def produced(self) -> int:
        return int(_h().omldm_ring_produced(self._ring))

    @property
    def consumed(self) -> int:
        return int(_h().omldm_ring_consumed(self._ring))

    def stop(self) -> None:
        if self._producer:
            _h().omldm_ring_close(self._ring)
            _h().omldm_producer_stop(self._producer)
            self._producer = None

    def __del__(self):
        try:
            self.stop()
            if self._ring:
                _h().omldm_ring_destroy(self._ring)
                self._ring = None
        except Exception:
            pass


class SparseTopic(Topic):
    """Ring of sparse mini-batches in fixed-width form (the "sparse" half of BASELINE config 5's
    sparse-to-dense stream).  A slot is one uint8 row: [int32 idx B*nnz | float val B*nnz | float y B]."""

    def __init__(self, d: int, batch: int, nnz: int, n_slots: int = 8, pin: Optional[bool] = None):
        self.d, self.batch, self.nnz, self.n_slots = d, batch, nnz, n_slots
        self.ld, self.wire = nnz, "f32"
        self.slot_bytes = batch * (2 * nnz + 1) * 4
        self.storage = torch.zeros(n_slots, self.slot_bytes, dtype=torch.uint8)
        if pin is None:
            pin = torch.cuda.is_available()
        if pin:
            self.storage = self.storage.pin_memory()
        self._ring = _h().omldm_ring_create(self.storage.data_ptr(), n_slots, self.slot_bytes)
        self._producer = None

    def start_synthetic(self, kind: int = KIND_SPARSE_MULTICLASS, seed: int = 0, n_aux: int = 10,
                        mode: int = MODE_GENERATE, max_batches: int = -1, n_threads: int = 1) -> None:
        super().start_synthetic(KIND_SPARSE_MULTICLASS, seed, n_aux, mode, max_batches, n_threads)

    @staticmethod
    def unpack(slot: torch.Tensor, batch: int, nnz: int):
        """(indices int32 [B*nnz], values fp32 [B*nnz], targets fp32 [B]) views of a (device) slot."""
        n = batch * nnz
        idx = slot[: 4 * n].view(torch.int32)
        val = slot[4 * n: 8 * n].view(torch.float32)
        tgt = slot[8 * n: 8 * n + 4 * batch].view(torch.float32)
        return idx, val, tgt


def generate_sparse_batch(d: int, batch: int, nnz: int, n_aux: int = 10, seed: int = 0, batch_index: int = 0,
                          pin: bool = False) -> torch.Tensor:
    out = torch.empty(batch * (2 * nnz + 1) * 4, dtype=torch.uint8)
    if pin and torch.cuda.is_available():
        out = out.pin_memory()
    _h().omldm_generate_sparse_batch(out.data_ptr(), d, nnz, batch, n_aux, seed, batch_index)
    return out


def generate_batch(kind: int, d: int, batch: int, seed: int = 0, batch_index: int = 0, n_aux: int = 0,
                   pin: bool = False, wire: str = "f32") -> torch.Tensor:
    """One synthetic staged batch [batch, pad32(d+2)] on the host (`wire="f16"`: compact wire rows, `"f16s"`: fp16
    staged rows)."""
    if wire == "f16s":
        out = torch.empty(batch, f16s_pitch(d), dtype=torch.float16)
        if pin and torch.cuda.is_available():
            out = out.pin_memory()
        _h().omldm_generate_batch_f16s(out.data_ptr(), kind, d, batch, n_aux, seed, batch_index)
        return out
    if wire == "f16":
        out = torch.empty(batch, wire_f16_pitch(d), dtype=torch.float16)
        if pin and torch.cuda.is_available():
            out = out.pin_memory()
        _h().omldm_generate_batch_f16(out.data_ptr(), kind, d, batch, n_aux, seed, batch_index)
        return out
    ld = pad32(d + 2)
    out = torch.empty(batch, ld, dtype=torch.float32)
    if pin and torch.cuda.is_available():
        out = out.pin_memory()
    _h().omldm_generate_batch(out.data_ptr(), kind, d, ld, batch, n_aux, seed, batch_index)
    return out


def parse_json_records(text: bytes, d: int, max_rows: int, n_threads: Optional[int] = None, offset: int = 0
                       ) -> Tuple[torch.Tensor, int, int]:
    """(staged rows [n, ld], n_forecast, bytes consumed); rows with valid flag 2.0 are forecasts.
    Parsing starts at byte `offset` of `text` (no copy of the remainder).  Large inputs are parsed on several
    threads (default: up to 8, one per 64 KB of text)."""
    ld = pad32(d + 2)
    out = torch.empty(max_rows, ld, dtype=torch.float32)
    nf = ctypes.c_int(0)
    used = ctypes.c_size_t(0)
    n_bytes = len(text) - offset
    if n_bytes <= 0:
        return out[:0], 0, 0
    if n_threads is None:
        import os
        n_threads = max(1, min(8, os.cpu_count() or 1, n_bytes >> 16))
    base = ctypes.cast(ctypes.c_char_p(text), ctypes.c_void_p).value      # `text` stays alive for the call
    n = _h().omldm_parse_json_records_mt(base + offset, n_bytes, d, out.data_ptr(), ld, max_rows, ctypes.byref(nf),
                                         ctypes.byref(used), n_threads)
    return out[:n], int(nf.value), int(used.value)
