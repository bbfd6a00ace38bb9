"""Compact fp16 wire format of the host->device link: native packers / producers, the staging op and the
engine's e2e path.  (The reference's wire format is JSON text, DataInstanceParser.scala:12-22.)"""
import ctypes

import numpy as np
import pytest
import torch

from omldm_b200.engine import Engine
from omldm_b200.ops import _lib
from omldm_b200.ops import functional as F
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters


def _bits(t):
    return t.contiguous().view(torch.int16)


def test_native_f32_to_f16_matches_ieee_round_to_nearest_even():
    h = _lib.host()
    h.omldm_pack_wire_f16.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                      ctypes.c_void_p]
    rng = np.random.default_rng(0)
    x = torch.from_numpy(rng.integers(0, 2 ** 32, size=1 << 18, dtype=np.uint64).astype(np.uint32).view(np.float32).copy())
    x = x[~torch.isnan(x)]
    special = torch.tensor([0.0, -0.0, 65504.0, 65519.9, 65520.0, 7e4, -7e4, float("inf"), -float("inf"), 5.96e-8,
                            2.98e-8, 2.9802322e-8, 2.9802326e-8, 6.1e-5, 6.0975e-5, 1.0, 1.0009765625, 1.00048828125,
                            1.0014648], dtype=torch.float32)
    x = torch.cat([x, special, torch.randn(50000), torch.randn(50000) * 1e-5, torch.randn(50000) * 1e-7, torch.ones(1)])
    d = x.numel() - 1
    out = torch.zeros(F.wire_f16_pitch(d), dtype=torch.float16)
    h.omldm_pack_wire_f16(x.data_ptr(), x.numel(), d, 1, d, out.data_ptr())
    assert torch.equal(_bits(out[:d]), _bits(x[:d].to(torch.float16)))
    assert float(out[-2:].view(torch.float32)) == 1.0          # the target keeps its fp32 bits


@pytest.mark.parametrize("d", [1, 5, 6, 28, 1000])
def test_pack_and_stage_roundtrip(d):
    raw = torch.randn(17, d + 1)
    Z = F.stage_rows(raw, d)
    w = F.pack_wire_f16(Z, d)
    assert w.shape == (17, F.wire_f16_pitch(d)) and F.wire_f16_pitch(d) % 2 == 0
    Z2 = F.stage_rows_f16(w, d)
    assert Z2.shape == Z.shape
    assert torch.equal(Z2[:, d], Z[:, d]) and torch.equal(Z2[:, d + 1], Z[:, d + 1])
    assert torch.equal(Z2[:, :d], Z[:, :d].to(torch.float16).float())
    assert float(Z2[:, d + 2:].abs().sum()) == 0.0
    assert torch.equal(_bits(F.pack_wire_f16(raw, d)), _bits(w))     # raw [B, d+1] rows pack the same way


@pytest.mark.parametrize("kind", [R.KIND_BINARY, R.KIND_REGRESSION, R.KIND_CLUSTERS, R.KIND_MULTICLASS])
def test_native_generator_f16_equals_packed_f32(kind):
    d = 11
    a = R.generate_batch(kind, d, 64, seed=5, batch_index=3, n_aux=4)
    h = R.generate_batch(kind, d, 64, seed=5, batch_index=3, n_aux=4, wire="f16")
    assert torch.equal(_bits(h), _bits(F.pack_wire_f16(a, d)))


def test_topic_f16_producer_and_put():
    d, B = 9, 32
    t = R.Topic(d, B, n_slots=4, wire="f16", pin=False)
    assert t.slot_bytes == B * F.wire_f16_pitch(d) * 2
    t.start_synthetic(R.KIND_REGRESSION, seed=1, max_batches=5)
    n = 0
    while True:
        g = t.get(2000)
        if g is None:
            break
        s, Zh, nr = g
        ref = R.generate_batch(R.KIND_REGRESSION, d, B, seed=1, batch_index=n, wire="f16")
        assert Zh.dtype == torch.float16 and nr == B and torch.equal(_bits(Zh), _bits(ref))
        t.release(s)
        n += 1
    assert n == 5
    t.stop()
    t2 = R.Topic(d, B, n_slots=2, wire="f16", pin=False)
    rows = R.generate_batch(R.KIND_BINARY, d, 20, seed=2)
    assert t2.put(rows)                          # staged fp32 rows are packed on the way in
    s, Zh, nr = t2.get(1000)
    assert nr == 20 and torch.equal(_bits(Zh[:20]), _bits(F.pack_wire_f16(rows, d)))
    with pytest.raises(ValueError):
        R.Topic(d, B, wire="int8")


def _engine(d, learner):
    e = Engine(JobParameters(parallelism=1, testSetSize=32), device="cpu")
    e.handle_request({"id": 1, "request": "Create", "learner": learner,
                      "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
    return e


def test_engine_e2e_path_accepts_wire_rows():
    d = 12
    a = _engine(d, {"name": "RegressorPA", "hyperParameters": {"C": 0.5}})
    b = _engine(d, {"name": "RegressorPA", "hyperParameters": {"C": 0.5}})
    for i in range(6):
        sa = a.train_batch_host(R.generate_batch(R.KIND_REGRESSION, d, 128, seed=3, batch_index=i), d)
        sb = b.train_batch_host(R.generate_batch(R.KIND_REGRESSION, d, 128, seed=3, batch_index=i, wire="f16"), d)
    wa, wb = a.pipelines[1].learner.w, b.pipelines[1].learner.w
    assert float((wa - wb).abs().max()) < 5e-3 * max(1.0, float(wa.abs().max()))
    assert int(sa[2]) == int(sb[2]) > 0                       # same number of records fitted
    assert abs(float(sa[0]) - float(sb[0])) < 1e-2 * max(1.0, abs(float(sa[0])))
    with pytest.raises(ValueError):
        b.train_batch_host(torch.zeros(4, 5, dtype=torch.float16), d)


@pytest.mark.gpu
def test_stage_rows_f16_kernel_matches_reference():
    """Kernel vs the plain torch restatement of the same op (exact: the conversion is a widening)."""
    for d, B in ((1, 3), (5, 7), (28, 1000), (1000, 4096), (999, 513)):
        raw = torch.randn(B, d + 1)
        w = F.pack_wire_f16(raw, d)
        ref = F.stage_rows_f16(w, d)                 # CPU tensors -> torch path
        got = F.stage_rows_f16(w.cuda(), d)          # CUDA tensors -> omldm_stage_rows_f16
        torch.cuda.synchronize()
        assert torch.equal(got.cpu(), ref), (d, B)
    # poisoned output buffer: every column of the staged row is written
    out = torch.full((64, F.pad32(30)), float("nan"), device="cuda")
    w = F.pack_wire_f16(torch.randn(64, 29), 28).cuda()
    F.stage_rows_f16(w, 28, out=out)
    assert not bool(torch.isnan(out).any())


@pytest.mark.gpu
def test_engine_e2e_wire_f16_gpu():
    d = 28
    eng = Engine(JobParameters(parallelism=1, testSetSize=32))      # same holdout size as the CPU engine of _engine()
    eng.handle_request({"id": 1, "request": "Create", "learner": {"name": "PA", "hyperParameters": {"C": 0.1}},
                        "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
    ref = _engine(d, {"name": "PA", "hyperParameters": {"C": 0.1}})
    for i in range(8):
        h = R.generate_batch(R.KIND_BINARY, d, 256, seed=4, batch_index=i, wire="f16", pin=True)
        st = eng.train_batch_host(h, d)
        ref.train_batch_host(h.clone(), d)
    torch.cuda.synchronize()
    assert int(st[2]) == ref.pipelines[1].learner.counters()["fitted"]
    assert torch.allclose(eng.pipelines[1].learner.w.cpu(), ref.pipelines[1].learner.w, rtol=1e-3, atol=1e-4)
    eng.close()
