"""Property tests (hypothesis) of the pure index / layout logic shared by host and device code."""
import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from omldm_b200.api.bucketing import bucket_response, reassemble
from omldm_b200.api.pojos import LearnerPOJO, QueryResponse
from omldm_b200.ops import functional as F
from omldm_b200.oracle.pipeline import HoldoutSampler
from omldm_b200.oracle.preprocessors import poly_factor_table
from omldm_b200.parallel.host_protocols import shard_bounds


@settings(max_examples=40, deadline=None)
@given(st.integers(1, 45000), st.integers(0, 3))
def test_bucketing_roundtrip(n, extra):
    params = {"a": [float(i) for i in range(n)], "b": 0.25}
    for k in range(extra):
        params[f"p{k}"] = list(range(k + 1))
    resp = QueryResponse(3, None, 1, [], LearnerPOJO("PA", {"C": 1}, params, None), "Synchronous", 5, .1, .2, .3)
    frags = bucket_response(resp)
    if len(frags) > 1:   # a response with at most two buckets is sent whole (FlinkNetwork.scala:179,187-191)
        assert all(len(v) <= 10000 for f in frags for v in (f.learner.parameters or {}).values() if isinstance(v, list))
    whole = reassemble(frags)
    assert whole.learner.parameters["a"] == params["a"] and whole.dataFitted == 5 and whole.protocol == "Synchronous"


@settings(max_examples=60, deadline=None)
@given(st.integers(1, 700), st.integers(0, 5000), st.integers(1, 40))
def test_holdout_index_map_equals_sequential_rule(B, count0, cap):
    """The in-place batch formulation (device kernel / torch op) == the reference's per-record rule."""
    sampler = HoldoutSampler(cap)
    sampler.count = count0
    h0 = F.holdouts_before(count0)
    for h in range(max(0, h0 - cap), h0):          # pre-fill the ring exactly as the stream would have
        sampler.test_set.append((np.array([-1.0 - h]), 0.0))
    ring = torch.zeros(cap, 4)
    for h in range(max(0, h0 - cap), h0):
        ring[h % cap, 0] = -1.0 - h
        ring[h % cap, 1] = 1.0
    Z = torch.zeros(B, 4)
    Z[:, 0] = torch.arange(B, dtype=torch.float32) + 1.0
    Z[:, 1] = 1.0
    expected = []
    for i in range(B):
        r = sampler.route(np.array([float(i + 1)]), 0.0)
        expected.append(None if r is None else float(r[0][0]))
    F.holdout_swap(Z, ring, 0, count0)
    got = [None if Z[i, 1] == 0 else float(Z[i, 0]) for i in range(B)]
    assert got == expected
    assert sorted(float(p[0][0]) for p in sampler.test_set) == sorted(float(v) for v in ring[ring[:, 1] != 0][:, 0])


@settings(max_examples=50, deadline=None)
@given(st.integers(0, 10000), st.integers(1, 16))
def test_shard_bounds_partition(n, H):
    b = shard_bounds(n, H)
    assert len(b) == H and b[0][0] == 0 and b[-1][1] == n
    assert all(b[i][1] == b[i + 1][0] for i in range(H - 1)) and all(lo <= hi for lo, hi in b)


@settings(max_examples=20, deadline=None)
@given(st.integers(1, 9), st.integers(1, 3))
def test_poly_table_counts_and_torch_op(d, degree):
    from math import comb
    tab = poly_factor_table(d, degree)
    assert len(tab) == sum(comb(d + k - 1, k) for k in range(1, degree + 1))
    x = torch.randn(5, d)
    Z = F.stage_rows(torch.cat([x, torch.zeros(5, 1)], 1), d)
    out = F.poly_features(Z, d, torch.from_numpy(tab), len(tab))
    ref = np.stack([np.prod(np.concatenate([xi, [1.0]])[tab], axis=1) for xi in x.numpy().astype(np.float64)])
    assert np.allclose(out[:, : len(tab)].numpy(), ref, rtol=1e-4, atol=1e-5)


@settings(max_examples=40, deadline=None)
@given(st.integers(1, 70), st.integers(1, 40), st.integers(0, 2 ** 31 - 1))
def test_wire_f16_roundtrip_properties(d, B, seed):
    """Compact wire rows: pitch is even and holds d features + an fp32 target; expanding them gives the staged row of the
    fp16-rounded features with the target's bits untouched, whatever d (odd d exercises the pad half)."""
    g = torch.Generator().manual_seed(seed)
    raw = torch.randn(B, d + 1, generator=g) * 10.0 ** float(torch.randint(-3, 4, (1,), generator=g))
    w = F.pack_wire_f16(raw, d)
    assert w.shape == (B, F.wire_f16_pitch(d)) and w.shape[1] % 2 == 0 and w.shape[1] - d in (2, 3)
    Z = F.stage_rows_f16(w, d)
    assert Z.shape == (B, F.pad32(d + 2))
    assert torch.equal(Z[:, :d], raw[:, :d].to(torch.float16).float())
    assert torch.equal(Z[:, d], torch.ones(B)) and torch.equal(Z[:, d + 1], raw[:, d])
    assert float(Z[:, d + 2:].abs().sum()) == 0.0


@settings(max_examples=40, deadline=None)
@given(st.integers(1, 12), st.integers(1, 30), st.integers(3, 200), st.integers(0, 2 ** 31 - 1))
def test_sparse_rows_dense_form_properties(nnz, B, d, seed):
    """Sparse rows -> dense rows: duplicate indices accumulate, out-of-range ones vanish, x.w computed from the sparse
    row equals the dense dot product."""
    g = torch.Generator().manual_seed(seed)
    idx = torch.randint(-2, d + 3, (B * nnz,), generator=g).int()       # some indices out of range on purpose
    val = torch.randn(B * nnz, generator=g)
    tgt = torch.randint(0, 4, (B,), generator=g).float()
    rows = F.sparse_rows(idx, val, tgt, B, nnz)
    assert rows.shape == (B, F.sparse_row_pitch(nnz)) and rows.shape[1] % 32 == 0
    Z = F.sparse_rows_to_dense(rows, nnz, d)
    w = torch.randn(d, generator=g)
    ok = (idx >= 0) & (idx < d)
    safe = torch.where(ok, idx, torch.zeros_like(idx)).long()
    sparse_dot = (torch.where(ok, val, torch.zeros_like(val)) * w[safe]).view(B, nnz).sum(1)
    assert torch.allclose(Z[:, :d] @ w, sparse_dot, rtol=1e-4, atol=1e-4)
    assert torch.equal(Z[:, d], torch.ones(B)) and torch.equal(Z[:, d + 1], tgt)


@settings(max_examples=30, deadline=None)
@given(st.integers(1, 9), st.lists(st.integers(0, 60), min_size=2, max_size=5), st.integers(0, 2 ** 31 - 1))
def test_scaler_moment_merge_equals_moments_of_the_union(d, sizes, seed):
    """Checkpoint rescaling: Chan's pairwise merge of (count, mean, M2) over several workers == one pass over all data."""
    from omldm_b200.utils.checkpoint import _merge_scaler
    rng = np.random.default_rng(seed)
    parts = [rng.standard_normal((n, d)) * 3.0 + 1.0 for n in sizes]
    states = []
    for X in parts:
        s = torch.zeros(1 + 2 * d, dtype=torch.float64)
        if len(X):
            s[0] = len(X)
            s[1:1 + d] = torch.from_numpy(X.mean(0))
            s[1 + d:] = torch.from_numpy(((X - X.mean(0)) ** 2).sum(0))
        states.append(s)
    merged = _merge_scaler(states)
    allx = np.concatenate(parts)
    assert float(merged[0]) == len(allx)
    if len(allx):
        assert np.allclose(merged[1:1 + d].numpy(), allx.mean(0), rtol=1e-9, atol=1e-9)
        assert np.allclose(merged[1 + d:].numpy(), ((allx - allx.mean(0)) ** 2).sum(0), rtol=1e-8, atol=1e-8)
