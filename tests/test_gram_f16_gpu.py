"""fp16 Gram kernel (csrc/gram_f16.cu: CTA pairs, tcgen05 cta_group::2, TMA, split-K, TMA reduce-add epilogue and the
in-kernel hub push) against a plain PyTorch reference of the same op: G = X^T X in fp64 from the same fp16 rows."""
import pytest
import torch

from omldm_b200.ops import functional as F

pytestmark = pytest.mark.gpu


def _rows(K, d, seed=0, empty_every=0):
    g = torch.Generator().manual_seed(seed)
    raw = torch.randn(K, d + 1, generator=g)
    raw[:, d] = raw[:, :d].sum(1) * 0.1 + 0.01 * torch.randn(K, generator=g)
    Zh = F.pack_f16s(raw, d)
    if empty_every:
        Zh[::empty_every] = 0          # empty rows (valid flag 0) contribute nothing
    return Zh.cuda()


def _ref(Zh, n):
    X = Zh[:, :n].double()
    return X.t() @ X, X.abs().t() @ X.abs()


def _upper_blocks(n, device):
    i = torch.arange(n, device=device)
    return (i[None, :] // 32 * 32 + 31) >= (i[:, None] // 32 * 32)


@pytest.mark.parametrize("K,d,ksplit", [(4096, 1000, 0), (300, 28, 0), (1000, 200, 1), (777, 509, 3), (64, 5, 0),
                                        (5000, 253, 0), (8192, 300, 5)])
def test_gram_matches_fp64_reference(K, d, ksplit):
    n = d + 3
    Zh = _rows(K, d, seed=K + d, empty_every=7)
    ldg = F.pad64(n)
    C = torch.zeros(ldg, ldg, device="cuda")
    F.gram_f16(Zh, n, C, ksplit=ksplit)
    F.gram_f16(Zh, n, C, ksplit=ksplit)                # accumulates
    torch.cuda.synchronize()
    ref, bound = _ref(Zh, n)
    keep = _upper_blocks(n, "cuda")
    got = C[:n, :n].double()
    err = (got - 2 * ref).abs()
    tol = 1e-5 * 2 * bound + 1e-5                      # exact fp16 products; the fp32 accumulation in TMEM truncates
    assert bool((err[keep] <= tol[keep]).all()), float((err / tol)[keep].max())
    assert float(got[~keep].abs().max()) == 0.0 if (~keep).any() else True     # blocks below the diagonal are not written
    if n < ldg:
        assert float(C[n:, :].abs().max()) == 0.0 and float(C[:, n:].abs().max()) == 0.0


@pytest.mark.parametrize("world,K,d", [(4, 2048, 1000), (2, 1500, 509), (8, 4096, 1000), (3, 640, 200)])
def test_hub_push_copies_rows_into_the_owners_slots(world, K, d):
    """Hub push with local buffers standing in for the peers: the split-K partials meet in the worker's running sum C,
    and the last CTA of every half tile copies the updated rows into the worker's slot at the owner of those rows -
    after every call the slots hold exactly the rows of C they own (a newer value replaces the older one) and nothing
    else."""
    n = d + 3
    ldg = F.pad64(n)
    rows_per_owner = ldg // world // 32 * 32
    if rows_per_owner * world < n:
        rows_per_owner += 32
    slots = [torch.zeros(ldg, ldg, device="cuda") for _ in range(world)]
    C = torch.zeros(ldg, ldg, device="cuda")
    scratch = F.gram_f16_scratch("cuda")
    peers = F.GramPeers()
    for p in range(world):
        peers.owner_base[p] = slots[p].data_ptr()
    peers.world, peers.rank, peers.rows_per_owner = world, 1 % world, rows_per_owner
    total = torch.zeros(n, n, dtype=torch.float64, device="cuda")
    bound = torch.zeros_like(total)
    keep = _upper_blocks(n, "cuda")
    for step in range(1, 4):
        Zh = _rows(K, d, seed=step)
        F.gram_f16(Zh, n, C, scratch=scratch, peers=peers)
        torch.cuda.synchronize()
        r, b = _ref(Zh, n)
        total += r
        bound += b
        tol = 1e-5 * bound + 1e-5
        err = (C[:n, :n].double() - total).abs()
        assert bool((err[keep] <= tol[keep]).all()), float((err / tol)[keep].max())
        for p in range(world):
            lo, hi = p * rows_per_owner, min(n, (p + 1) * rows_per_owner)
            got = slots[p][:n, :n]
            if lo < hi:
                assert torch.equal(got[lo:hi][keep[lo:hi]], C[:n, :n][lo:hi][keep[lo:hi]]), (step, p)
            mask = torch.ones(n, dtype=torch.bool, device="cuda")
            mask[lo:hi] = False
            if mask.any():
                assert float(got[mask].abs().max()) == 0.0, f"owner {p} received foreign rows"


@pytest.mark.parametrize("world,K,d", [(4, 16384, 1000), (2, 1500, 509), (8, 32768, 1000), (3, 640, 200)])
def test_hub_push_lead_mode_ships_the_previous_result(world, K, d):
    """Lead mode: while the MMAs of call t run, the kernel copies the running sum as it was before the call.  With a main
    loop much longer than the copy (the bench shape) the slots hold exactly the result of call t - 1 after call t; with
    short loops a copy may overlap another CTA's epilogue, which is why readers always run the catch-up push first
    (parallel/fused.py::ShardedGramDriver.gather) - there only the placement and the running sum are checked."""
    exact = K >= 16384 and d >= 1000
    n = d + 3
    ldg = F.pad64(n)
    rows_per_owner = ldg // world // 32 * 32
    if rows_per_owner * world < n:
        rows_per_owner += 32
    slots = [torch.zeros(ldg, ldg, device="cuda") for _ in range(world)]
    C = torch.zeros(ldg, ldg, device="cuda")
    scratch = F.gram_f16_scratch("cuda")
    peers = F.GramPeers()
    for p in range(world):
        peers.owner_base[p] = slots[p].data_ptr()
    peers.world, peers.rank, peers.rows_per_owner, peers.lead = world, 1 % world, rows_per_owner, 1
    keep = _upper_blocks(n, "cuda")
    before = C.clone()
    for step in range(1, 4):
        Zh = _rows(K, d, seed=step)
        F.gram_f16(Zh, n, C, scratch=scratch, peers=peers)
        torch.cuda.synchronize()
        for p in range(world):
            lo, hi = p * rows_per_owner, min(n, (p + 1) * rows_per_owner)
            got = slots[p][:n, :n]
            if lo < hi and exact:
                assert torch.equal(got[lo:hi][keep[lo:hi]], before[:n, :n][lo:hi][keep[lo:hi]]), (step, p)
            mask = torch.ones(n, dtype=torch.bool, device="cuda")
            mask[lo:hi] = False
            if mask.any():
                assert float(got[mask].abs().max()) == 0.0, f"owner {p} received foreign rows"
        before = C.clone()
    # same running sum as without any push
    C2 = torch.zeros(ldg, ldg, device="cuda")
    for step in range(1, 4):
        F.gram_f16(_rows(K, d, seed=step), n, C2)
    torch.cuda.synchronize()
    assert torch.allclose(C, C2, rtol=1e-5, atol=1e-3)


def test_gram_f16_speed_report():
    """Not an assertion on speed: prints the device time at the bench shape so a GPU run leaves the number behind."""
    K, d = 32768, 1000
    n = d + 3
    pool = [_rows(K, d, seed=s) for s in range(3)]          # 3 x 64 MB > L2 in total
    C = torch.zeros(1024, 1024, device="cuda")
    for Zh in pool:
        F.gram_f16(Zh, n, C)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    reps = 30
    for i in range(reps):
        F.gram_f16(pool[i % 3], n, C)
    e1.record()
    torch.cuda.synchronize()
    us = 1000.0 * e0.elapsed_time(e1) / reps
    flops = 2.0 * K * 10 * 256 * 256            # 10 upper 256 x 256 tiles
    print(f"\ngram_f16 K={K} d={d}: {us:.1f} us per call, {flops / us / 1e6:.0f} TFLOP/s executed")
    for line in gram_phase_report():
        print(line)


def gram_phase_report():
    """Per-CTA phase durations of the LAST gram_f16 launch; empty unless the loaded library is a tuning build
    (scripts/build_variant.sh timing csrc/gram_f16.cu -DOMLDM_GRAM_TIMING)."""
    from omldm_b200.ops import _lib
    lib = _lib.kernels()
    if not hasattr(lib, "omldm_gram_f16_stamps"):
        return []
    import ctypes
    import numpy as np
    torch.cuda.synchronize()
    buf = np.zeros(160 * 16, dtype=np.int64)
    assert lib.omldm_gram_f16_stamps(buf.ctypes.data_as(ctypes.c_void_p)) == 0
    st = buf.reshape(160, 16)[:140]
    out = []
    names = ["prologue", "first stage landed", "main loop", "epilogue issue", "reduce drain", "teardown"]
    for i, nm in enumerate(names):
        dt = st[:, i + 1] - st[:, i]
        ok = (st[:, i + 1] > 0) & (st[:, i] > 0)
        if ok.any():
            out.append(f"  {nm:20s} cycles: median {int(np.median(dt[ok]))}  min {int(dt[ok].min())}  max {int(dt[ok].max())}")
    full = buf.reshape(160, 16)[:140]
    for rk in (0, 1):                                   # loss warps of the diagonal-tile CTAs, by rank in the pair
        sel = (full[:, 11] > 0) & ((np.arange(140) % 20) % 2 == rk)
        if sel.any():
            out.append(f"  loss warp, rank {rk}: waiting {int(np.median(full[sel, 10]))} cycles, working {int(np.median(full[sel, 11]))} cycles (median of {int(sel.sum())} CTAs)")
    diag_tiles = {0, 4, 7, 9}
    for name, sel_fn in (("diagonal", lambda t: t in diag_tiles), ("off-diagonal", lambda t: t not in diag_tiles)):
        idx = np.array([i for i in range(140) if sel_fn((i % 20) // 2)])
        lead, both = idx[idx % 2 == 0], idx
        if full[both, 12].any():
            out.append(f"  {name} tiles: producer waited for a free stage {int(np.median(full[both, 12]))} cycles, "
                       f"MMA warp waited for data {int(np.median(full[lead, 13]))} cycles (totals over the main loop, median)")
    ok = st[:, 6] > 0
    if ok.any():
        tot, ns = (st[:, 6] - st[:, 0])[ok], (st[:, 9] - st[:, 8])[ok]
        out.append(f"  leader CTAs start..teardown: median {int(np.median(tot))} cycles = {np.median(ns) / 1e3:.1f} us "
                   f"-> {np.median(tot / np.maximum(ns, 1)):.3f} GHz; first start .. last end {(st[:, 9][ok].max() - st[:, 8][st[:, 8] > 0].min()) / 1e3:.1f} us")
    return out


@pytest.mark.parametrize("B,d", [(4096, 1000), (300, 28), (5000, 253), (32768, 1000)])
def test_loss_pass_in_gram_order_matches_reference(B, d):
    """sq_residual_f16s with the rows walked in the order of the Gram kernel's batch-axis splits (the layout used when the
    two run side by side) gives the same sums as the plain pass and as PyTorch."""
    Zh = _rows(B, d, seed=B, empty_every=9)
    wt = torch.randn(d + 3, device="cuda") * 0.1
    wt[d + 1:] = -1.0
    valid = Zh[:, d] != 0
    r = (Zh[:, : d + 3].double() @ wt.double())[valid]
    ref = float((r ** 2).sum())
    for like_gram in (False, True):
        st = torch.zeros(4, dtype=torch.float64, device="cuda")
        F.sq_residual_f16s(Zh, d, wt, st, like_gram=like_gram)
        torch.cuda.synchronize()
        assert int(st[2]) == int(valid.sum())
        assert abs(float(st[0]) - ref) <= 1e-5 * max(1.0, ref), (like_gram, float(st[0]), ref)


@pytest.mark.parametrize("K,d,empty_every", [(32768, 1000, 0), (4096, 1000, 7), (1500, 509, 5), (640, 200, 3), (300, 28, 4),
                                             (20000, 125, 0)])
def test_loss_inside_the_gram_kernel_matches_the_separate_pass(K, d, empty_every):
    """`gram_f16(..., loss=)`: the diagonal-tile CTAs form the residual dots from the shared-memory stages while the MMAs
    run.  Same squared-residual sum and valid-row count as sq_residual_f16s and as PyTorch in fp64, the same Gram matrix
    as without it, the learning-curve sample written, and the workspace left zero - three calls in a row."""
    n = d + 3
    ldg = F.pad64(n)
    C = torch.zeros(ldg, ldg, device="cuda")
    C2 = torch.zeros(ldg, ldg, device="cuda")
    wt = (torch.randn(n, device="cuda") * 0.1)
    wt[d + 1:] = -1.0
    st = torch.zeros(4, dtype=torch.float64, device="cuda")
    st_ref = torch.zeros(4, dtype=torch.float64, device="cuda")
    dots, words = F.gram_loss_workspace(K, "cuda")
    ring = torch.zeros(8, 2, dtype=torch.float64, device="cuda")
    cnt = torch.zeros(2, dtype=torch.int64, device="cuda")
    desc = F.make_gram_loss(wt, d, st, dots, words, ring, cnt)
    exact = 0.0
    nvalid = 0
    for step in range(3):
        Zh = _rows(K, d, seed=10 + step, empty_every=empty_every)
        F.gram_f16(Zh, n, C, loss=desc)
        F.gram_f16(Zh, n, C2)
        F.sq_residual_f16s(Zh, d, wt, st_ref)
        torch.cuda.synchronize()
        valid = Zh[:, d] != 0
        r = (Zh[:, :n].double() @ wt.double())[valid]
        exact += float((r * r).sum())
        nvalid += int(valid.sum())
        assert float(st[2]) == nvalid == float(st_ref[2]), (step, float(st[2]), nvalid)
        assert abs(float(st[0]) - exact) <= 1e-4 * max(exact, 1.0), (step, float(st[0]), exact)
        assert abs(float(st[0]) - float(st_ref[0])) <= 1e-4 * max(exact, 1.0)
        assert int(cnt[0]) == step + 1 and float(ring[step, 0]) == float(st[0]) and float(ring[step, 1]) == nvalid
        assert float(dots.abs().max()) == 0.0 and int(words.abs().max()) == 0          # workspace clean for the next call
    assert torch.allclose(C, C2, rtol=1e-5, atol=1e-3)         # split-K partials meet in a different order
