"""MultiClassPA straight from sparse rows (csrc/sparse_mcpa.cu, Engine `sparseDirect`): same model, statistics,
holdout and responses as the sparse->dense route, without the dense batch."""
import json

import pytest
import torch

from omldm_b200.engine import Engine
from omldm_b200.ops import functional as F
from omldm_b200.stream import ring as R
from omldm_b200.utils.checkpoint import load_checkpoint, save_checkpoint
from omldm_b200.utils.config import JobParameters

D, B, NNZ, C = 5000, 64, 16, 6


def _engine(direct, device="cpu", d=D, tc=None):
    jp = JobParameters(parallelism=1, testSetSize=16)
    jp.extra["sparseDirect"] = "true" if direct else "false"
    e = Engine(jp, device=device)
    e.handle_request({"id": 1, "request": "Create",
                      "learner": {"name": "MultiClassPA", "hyperParameters": {"nClasses": C, "C": 0.1}},
                      "trainingConfiguration": dict({"protocol": "Synchronous", "nFeatures": d}, **(tc or {}))})
    return e


def _slot(i, d=D, unique=True):
    """One ELL slot; unique column indices per record make the two routes agree exactly (with duplicates the
    dense route's step size sees (sum v)^2, the sparse one sum v^2)."""
    g = torch.Generator().manual_seed(100 + i)
    if unique:
        idx = torch.stack([torch.randperm(d, generator=g)[:NNZ] for _ in range(B)]).int()
    else:
        idx = torch.randint(0, d, (B, NNZ), generator=g).int()
    val = torch.randn(B, NNZ, generator=g)
    y = torch.randint(0, C, (B,), generator=g).float()
    return torch.cat([idx.reshape(-1).view(torch.uint8), val.reshape(-1).view(torch.uint8), y.view(torch.uint8)])


def test_sparse_rows_layout_and_dense_form():
    slot = R.generate_sparse_batch(D, B, NNZ, n_aux=C, seed=1, batch_index=0)
    idx, val, tgt = R.SparseTopic.unpack(slot, B, NNZ)
    rows = F.sparse_rows(idx, val, tgt, B, NNZ)
    assert rows.shape == (B, F.sparse_row_pitch(NNZ)) and rows.shape[1] % 32 == 0
    assert torch.equal(rows[:, :NNZ].contiguous().view(torch.int32), idx.view(B, NNZ))
    assert torch.equal(rows[:, NNZ:2 * NNZ], val.view(B, NNZ))
    assert bool((rows[:, 2 * NNZ] == 1).all()) and torch.equal(rows[:, 2 * NNZ + 1], tgt)
    ip = torch.arange(0, (B + 1) * NNZ, NNZ, dtype=torch.int32)
    Zref, _ = F.densify(ip, idx, val, tgt, D)
    assert torch.equal(F.sparse_rows_to_dense(rows, NNZ, D), Zref)
    rows[3] = 0.0                                    # an empty row stays empty
    assert float(F.sparse_rows_to_dense(rows, NNZ, D)[3].abs().sum()) == 0.0


def test_direct_route_equals_dense_route():
    a, b = _engine(False), _engine(True)
    for i in range(12):
        s = _slot(i)
        a.train_sparse(s.clone(), B, NNZ, D)
        b.train_sparse(s.clone(), B, NNZ, D)
    la, lb = a.pipelines[1].learner, b.pipelines[1].learner
    assert torch.allclose(la.W, lb.W, rtol=1e-5, atol=1e-6)
    ca, cb = la.counters(), lb.counters()
    assert ca["fitted"] == cb["fitted"] and ca["aux"] == cb["aux"]
    assert abs(ca["loss_sum"] - cb["loss_sum"]) < 1e-6 * ca["loss_sum"]
    assert b.holdout.shape == (16, F.sparse_row_pitch(NNZ)) and a.holdout.shape[1] == F.pad32(D + 2)
    ra = json.loads(a.handle_request({"id": 1, "request": "Query", "requestId": 5})[-1])
    rb = json.loads(b.handle_request({"id": 1, "request": "Query", "requestId": 5})[-1])
    assert ra["dataFitted"] == rb["dataFitted"] and ra["score"] == rb["score"]
    assert abs(ra["loss"] - rb["loss"]) < 1e-6


def test_direct_route_guards():
    b = _engine(True)
    b.train_sparse(_slot(0), B, NNZ, D)
    with pytest.raises(ValueError):                   # dense batches cannot be mixed in afterwards
        b.train_batch(torch.zeros(4, F.pad32(D + 2)), D)
    # a pipeline with a preprocessor cannot take sparse rows: the engine silently uses the dense route
    jp = JobParameters(parallelism=1, testSetSize=16)
    jp.extra["sparseDirect"] = "true"
    e = Engine(jp, device="cpu")
    e.handle_request({"id": 1, "request": "Create", "learner": {"name": "MultiClassPA", "hyperParameters": {"nClasses": C}},
                      "preProcessors": [{"name": "MinMaxScaler"}],
                      "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": 200}})
    e.train_sparse(_slot(0, d=200), B, NNZ, 200)
    assert e._sparse_nnz is None and e.holdout.shape[1] == F.pad32(202)


def test_direct_route_checkpoint_roundtrip(tmp_path):
    a = _engine(True)
    for i in range(4):
        a.train_sparse(_slot(i), B, NNZ, D)
    save_checkpoint(a, str(tmp_path), step=4)
    b = _engine(True)
    assert load_checkpoint(b, str(tmp_path)) == 4
    assert b._sparse_nnz == NNZ
    for i in range(4, 8):
        a.train_sparse(_slot(i), B, NNZ, D)
        b.train_sparse(_slot(i), B, NNZ, D)
    assert torch.allclose(a.pipelines[1].learner.W, b.pipelines[1].learner.W, atol=1e-6)
    assert torch.equal(a.holdout, b.holdout)


@pytest.mark.gpu
def test_sparse_kernels_match_reference_route():
    """GPU sparse kernels vs the torch restatement (CPU engine, same sparse semantics incl. duplicates)."""
    ref, gpu = _engine(True, "cpu"), _engine(True, "cuda")
    for i in range(10):
        s = _slot(i, unique=(i % 2 == 0))
        ref.train_sparse(s.clone(), B, NNZ, D)
        gpu.train_sparse(s.cuda(), B, NNZ, D)
    torch.cuda.synchronize()
    lr, lg = ref.pipelines[1].learner, gpu.pipelines[1].learner
    assert torch.allclose(lr.W, lg.W.cpu(), rtol=1e-4, atol=1e-5)
    cr, cg = lr.counters(), lg.counters()
    assert cr["fitted"] == cg["fitted"] and abs(cr["aux"] - cg["aux"]) <= 2
    assert abs(cr["loss_sum"] - cg["loss_sum"]) < 1e-4 * cr["loss_sum"]
    assert torch.equal(ref.holdout, gpu.holdout.cpu())
    gpu.close()


@pytest.mark.gpu
def test_sparse_rows_kernel():
    slot = R.generate_sparse_batch(D, B, NNZ, n_aux=C, seed=3, batch_index=1)
    idx, val, tgt = R.SparseTopic.unpack(slot, B, NNZ)
    ref = F.sparse_rows(idx, val, tgt, B, NNZ)
    out = torch.full((B, F.sparse_row_pitch(NNZ)), float("nan"), device="cuda")
    idc, vc, tc = R.SparseTopic.unpack(slot.cuda(), B, NNZ)
    F.sparse_rows(idc, vc, tc, B, NNZ, out=out)
    torch.cuda.synchronize()
    assert torch.equal(out.cpu().view(torch.int32), ref.view(torch.int32))
