"""Protocol state machines of the fused kernels (csrc/sync.cu), exercised on CPU through the fake
symmetric-memory backend: one thread per rank, real interleavings, no GPU.

The reference tests none of this (no test tree); the invariants are the ones its protocols
promise: Synchronous = every replica equals the average after a step, GM = nothing moves unless
somebody left the ball, Asynchronous = the hub holds every push, SSP = bounded staleness, EASGD =
elastic moves conserve mass, FGM = replicas agree after every round and rounds do happen.
"""
import math

import numpy as np
import pytest

from omldm_b200.parallel.fake_backend import PAD_CLOCK, PAD_FGM, PAD_STATUS, FakeWorld


def _status(w):
    return [int(p[PAD_STATUS]) for p in w.pads]


@pytest.mark.parametrize("k", [2, 3, 8])
@pytest.mark.parametrize("n", [5, 64, 1000])
def test_allreduce_mean_repeated(k, n):
    w = FakeWorld(k, n)
    rng = np.random.default_rng(k * 1000 + n)
    for it in range(4):       # flags are monotonic: back-to-back epochs must not need a reset
        xs = [rng.standard_normal(n) for _ in range(k)]
        for r in range(k):
            w.bufs[r][:] = xs[r]
        w.run(lambda rk: rk.allreduce(0))
        ref = np.mean(xs, axis=0)
        for r in range(k):
            np.testing.assert_allclose(w.bufs[r], ref, rtol=1e-12, atol=1e-12)
            np.testing.assert_allclose(w.ranks[r].e, ref, rtol=1e-12, atol=1e-12)
    assert _status(w) == [0] * k


def test_allreduce_sum_delta():
    k, n = 4, 257
    w = FakeWorld(k, n)
    rng = np.random.default_rng(1)
    g0 = rng.standard_normal(n)
    deltas = [rng.standard_normal(n) for _ in range(k)]
    for r in range(k):
        w.ranks[r].e[:] = g0
        w.bufs[r][:] = deltas[r]
    w.run(lambda rk: rk.allreduce(1))
    for r in range(k):
        np.testing.assert_allclose(w.ranks[r].e, g0 + np.sum(deltas, axis=0), rtol=1e-12, atol=1e-12)
        assert not w.bufs[r].any()


def test_gm_votes():
    k, n = 4, 64
    w = FakeWorld(k, n)
    for r in range(k):
        w.bufs[r][:] = 0.001 * (r + 1)
    before = [b.copy() for b in w.bufs]
    out = []
    w.run(lambda rk: out.append(rk.gm(10.0)))
    assert out == [False] * k
    for r in range(k):
        np.testing.assert_array_equal(w.bufs[r], before[r])
    w.bufs[2][:] = 1.0           # rank 2 leaves the ball: everybody must see the violation
    ref = np.mean([b.copy() for b in w.bufs], axis=0)
    out.clear()
    w.run(lambda rk: out.append(rk.gm(10.0)))
    assert out == [True] * k
    for r in range(k):
        np.testing.assert_allclose(w.bufs[r], ref)
        np.testing.assert_allclose(w.ranks[r].e, ref)
    # a quiet step right after the sync: votes of the previous epoch must not leak into this one
    out.clear()
    w.run(lambda rk: out.append(rk.gm(10.0)))
    assert out == [False] * k
    assert _status(w) == [0] * k


def test_async_hub_collects_every_push():
    k, n = 4, 100
    w = FakeWorld(k, n)
    rng = np.random.default_rng(2)
    w0 = rng.standard_normal(n)
    upd = [rng.standard_normal(n) for _ in range(k)]
    for r in range(k):
        w.bufs[r][:] = w0 + upd[r]
        w.hubs[r][:] = w0
        w.ranks[r].e[:] = w0
    w.run(lambda rk: rk.async_push_pull(0))
    ref = w0 + np.mean(upd, axis=0)
    for r in range(k):
        lo, hi = w.shard(r)
        np.testing.assert_allclose(w.hubs[r][lo:hi], ref[lo:hi], rtol=1e-12, atol=1e-12)
        np.testing.assert_array_equal(w.bufs[r], w.ranks[r].e)     # worker adopts what it pulled


def test_async_additive_mode():
    k, n = 3, 40
    w = FakeWorld(k, n)
    rng = np.random.default_rng(3)
    deltas = [rng.standard_normal(n) for _ in range(k)]
    for r in range(k):
        w.bufs[r][:] = deltas[r]
    w.run(lambda rk: rk.async_push_pull(1))
    tot = np.sum(deltas, axis=0)
    for r in range(k):
        lo, hi = w.shard(r)
        np.testing.assert_allclose(w.hubs[r][lo:hi], tot[lo:hi], rtol=1e-12, atol=1e-12)
        assert not w.bufs[r].any()
    # the last pusher of each shard saw the full sum; nobody saw more than it
    for r in range(k):
        lo, hi = w.shard(r)
        seen = np.stack([w.ranks[q].e[lo:hi] for q in range(k)])
        assert any(np.allclose(seen[q], tot[lo:hi]) for q in range(k))


@pytest.mark.parametrize("staleness", [0, 1, 3])
def test_ssp_bounded_staleness(staleness):
    k, n, steps = 4, 16, 12
    w = FakeWorld(k, n)
    worst = [0] * k

    def body(rk):
        import time
        for c in range(steps):
            if rk.me == 0:
                time.sleep(0.002)          # a straggler: the others must wait for it
            rk.buf[:] += 1.0
            rk.ssp(staleness)
            lag = rk.clock - min(int(w.pads[rk.me][PAD_CLOCK + t]) for t in range(k))
            worst[rk.me] = max(worst[rk.me], lag)
    w.run(body)
    assert max(worst) <= staleness
    assert _status(w) == [0] * k
    for r in range(k):
        assert [int(v) for v in w.pads[r][PAD_CLOCK:PAD_CLOCK + k]] == [steps] * k


def test_ssp_timeout_sets_status_instead_of_hanging():
    w = FakeWorld(2, 8, spin_timeout_s=0.05)
    w.ranks[0].ssp(0)              # rank 1 never arrives
    assert int(w.pads[0][PAD_STATUS]) == 2


def test_barrier_timeout_sets_status():
    w = FakeWorld(2, 8, spin_timeout_s=0.05)
    w.ranks[0].barrier()
    assert int(w.pads[0][PAD_STATUS]) == 1


def test_easgd_conserves_mass_and_contracts():
    k, n = 4, 50
    w = FakeWorld(k, n)
    rng = np.random.default_rng(4)
    centre = rng.standard_normal(n)
    xs = [centre + rng.standard_normal(n) for _ in range(k)]
    for r in range(k):
        w.bufs[r][:] = xs[r]
        w.hubs[r][:] = centre
    w.run(lambda rk: rk.easgd(0.25))
    new_centre = np.concatenate([w.hubs[r][slice(*w.shard(r))] for r in range(k)])
    np.testing.assert_allclose(sum(xs) + centre, sum(w.bufs) + new_centre, rtol=1e-10, atol=1e-10)
    assert sum(np.linalg.norm(b - new_centre) for b in w.bufs) < sum(np.linalg.norm(x - centre) for x in xs)


@pytest.mark.parametrize("k", [2, 4, 8])
def test_fgm_rounds_and_agreement(k):
    n, T = 32, 1.0
    w = FakeWorld(k, n)
    w.ranks[0].fgm_init_hub(T)
    rounds_seen = 0
    for step in range(200):
        for r in range(k):
            w.bufs[r][:] += 0.002 * (r + 1)
        done = []
        w.run(lambda rk: done.append(rk.fgm_step(T)))
        assert len(set(done)) == 1, "ranks disagree on whether this step ended a round"
        if done[0]:
            rounds_seen += 1
            for r in range(1, k):
                np.testing.assert_allclose(w.bufs[r], w.bufs[0], rtol=1e-12, atol=1e-12)
                np.testing.assert_allclose(w.ranks[r].e, w.bufs[r], rtol=1e-12, atol=1e-12)
        avg = np.mean([w.bufs[r] - w.ranks[r].e for r in range(k)], axis=0)
        assert np.linalg.norm(avg) <= math.sqrt(T) * 1.5
    assert rounds_seen >= 3
    assert all(rk.fgm["rounds"] == rounds_seen for rk in w.ranks)
    assert _status(w) == [0] * k


def test_fgm_quiet_workers_never_synchronise():
    k, n, T = 4, 16, 1.0
    w = FakeWorld(k, n)
    w.ranks[0].fgm_init_hub(T)
    for step in range(50):
        for r in range(k):
            w.bufs[r][:] = 1e-4 * math.sin(step + r)     # drift stays far inside the safe zone
        done = []
        w.run(lambda rk: done.append(rk.fgm_step(T)))
        assert done == [False] * k
    assert int(w.pads[0][PAD_FGM + 3]) == 0


def test_fgm_flush_completes_a_pending_round():
    """A round flagged by the hub during the LAST step of a run is pending on every rank until one
    more kernel runs - FGMDriver.flush()/Engine.flush() exist for exactly this."""
    k, n, T = 4, 16, 1.0
    w = FakeWorld(k, n)
    w.ranks[0].fgm_init_hub(T)
    pending = False
    for step in range(400):
        for r in range(k):
            w.bufs[r][:] += 0.01
        w.run(lambda rk: rk.fgm_step(T))
        if any(rk.fgm_pending() for rk in w.ranks):
            pending = True
            break
    assert pending, "drifting workers must trigger a round"
    assert all(rk.fgm_pending() for rk in w.ranks)       # the flag is written to every pad
    w.run(lambda rk: rk.fgm_step(T))                     # the flush step
    assert not any(rk.fgm_pending() for rk in w.ranks)
    for r in range(1, k):
        np.testing.assert_allclose(w.bufs[r], w.bufs[0])


def test_barrier_and_query_reduce():
    k = 8
    w = FakeWorld(k, 4)
    out = [None] * k

    def body(rk):
        rk.barrier()
        out[rk.me] = rk.query_reduce(np.arange(4.0) + 4 * rk.me)
        rk.barrier()
    w.run(body)
    ref = np.arange(4 * k, dtype=float).reshape(k, 4).sum(0)
    for r in range(k):
        np.testing.assert_allclose(out[r], ref)
    assert _status(w) == [0] * k


def test_mixed_sequence_shares_the_flag_space():
    """allreduce, barrier, query_reduce all use the ENTRY/EXIT flags: sequences stay monotonic."""
    k, n = 4, 33
    w = FakeWorld(k, n)

    def body(rk):
        for it in range(5):
            rk.buf[:] = rk.me + it
            rk.allreduce(0)
            rk.barrier()
            s = rk.query_reduce(np.full(4, float(rk.me)))
            assert s[0] == sum(range(k))
            assert np.allclose(rk.buf, np.mean(range(k)) + it)
    w.run(body)
    assert _status(w) == [0] * k


# ---------------------------------------------------------------------------------------------------------
# The same invariants under randomised interleavings: every rank pauses for a random time wherever the kernels
# touch peer memory (FakeWorld(jitter=...)), several seeds each.
@pytest.mark.parametrize("seed", range(6))
def test_fgm_lockstep_decisions_under_jitter(seed):
    k, n, T = 4, 16, 1.0
    w = FakeWorld(k, n, jitter=2e-4, seed=seed)
    w.ranks[0].fgm_init_hub(T)
    rounds = 0
    for step in range(120):
        for r in range(k):
            w.bufs[r][:] += 0.004 * (r + 1)
        done = []
        w.run(lambda rk: done.append(rk.fgm_step(T)))
        assert len(set(done)) == 1, f"step {step}: ranks took different branches {done}"
        if done[0]:
            rounds += 1
            for r in range(1, k):
                np.testing.assert_allclose(w.bufs[r], w.bufs[0], rtol=1e-12, atol=1e-12)
    assert rounds >= 2 and _status(w) == [0] * k


@pytest.mark.parametrize("seed", range(6))
def test_free_running_fgm_with_drain_flush_under_jitter(seed):
    """Ranks step without any per-step rendezvous (as separate GPUs do), with DIFFERENT numbers of steps; the
    drain flush (done flags at the hub, hub-issued terminate word) leaves nobody inside a round-end barrier,
    whatever the skew - including a rank that finishes long before the hub announces a round."""
    k, n, T = 4, 16, 1.0
    steps = [60, 25, 60, 48]
    w = FakeWorld(k, n, jitter=3e-4, seed=seed, spin_timeout_s=30.0)
    w.ranks[0].fgm_init_hub(T)

    def body(rk):
        for s in range(steps[rk.me]):
            rk.buf[:] += 0.01 * (rk.me + 1)
            rk.fgm_step(T)
        rk.fgm_flush(T, gen=1)
    w.run(body)
    assert _status(w) == [0] * k
    assert not any(rk.fgm_pending() for rk in w.ranks)
    assert len({rk.fgm["rounds"] for rk in w.ranks}) == 1 and w.ranks[0].fgm["rounds"] >= 1
    # a second stream segment and a second flush generation on the same pads
    def body2(rk):
        for s in range(10 + 5 * rk.me):
            rk.buf[:] += 0.02
            rk.fgm_step(T)
        rk.fgm_flush(T, gen=2)
    w.run(body2)
    assert _status(w) == [0] * k and not any(rk.fgm_pending() for rk in w.ranks)
    assert len({rk.fgm["rounds"] for rk in w.ranks}) == 1


def test_drain_flush_in_lockstep_emulation():
    """One launch for all ranks per step (emulate mode): the flush loop ends in the same launch on every rank."""
    k, n, T = 4, 16, 1.0
    w = FakeWorld(k, n, jitter=1e-4, seed=7)
    w.ranks[0].fgm_init_hub(T)
    for step in range(80):
        for r in range(k):
            w.bufs[r][:] += 0.01
        w.run(lambda rk: rk.fgm_step(T))
    for launch in range(20):
        w.run(lambda rk: rk.fgm_step(T, drain=1))
        term = [int(w.pads[r][PAD_FGM + 5]) == 1 for r in range(k)]
        assert len(set(term)) == 1
        if term[0]:
            break
    assert term[0] and launch <= 3
    assert _status(w) == [0] * k and not any(rk.fgm_pending() for rk in w.ranks)


@pytest.mark.parametrize("seed", range(4))
def test_gm_and_allreduce_under_jitter(seed):
    k, n = 5, 37
    w = FakeWorld(k, n, jitter=2e-4, seed=seed)
    rng = np.random.default_rng(seed)
    for it in range(10):
        xs = [rng.standard_normal(n) * (3.0 if it % 3 == 0 else 0.01) for _ in range(k)]
        for r in range(k):
            w.bufs[r][:] = w.ranks[r].e + xs[r]
        expect_sync = any(float(x @ x) > 0.5 for x in xs)
        ref = np.mean([b.copy() for b in w.bufs], axis=0)
        out = []
        w.run(lambda rk: out.append(rk.gm(0.5)))
        assert out == [expect_sync] * k
        if expect_sync:
            for r in range(k):
                np.testing.assert_allclose(w.bufs[r], ref, rtol=1e-12, atol=1e-12)
    assert _status(w) == [0] * k


@pytest.mark.parametrize("seed", range(3))
def test_async_conservation_under_jitter(seed):
    """Whatever the interleaving of pushes, each hub shard ends up with the initial model plus the mean of all
    pushed updates (additive atomics commute)."""
    k, n, steps = 4, 24, 5
    w = FakeWorld(k, n, jitter=2e-4, seed=seed)
    rng = np.random.default_rng(10 + seed)
    upd = rng.standard_normal((k, steps, n))

    def body(rk):
        for s in range(steps):
            rk.buf[:] += upd[rk.me, s]
            rk.async_push_pull(0)
    w.run(body)
    ref = upd.sum((0, 1)) / k
    for r in range(k):
        lo, hi = w.shard(r)
        np.testing.assert_allclose(w.hubs[r][lo:hi], ref[lo:hi], rtol=1e-10, atol=1e-10)


@pytest.mark.parametrize("seed", range(5))
def test_design_model_fused_reduce_scatter_three_buffers(seed):
    """Protocol model of the planned GEMM + reduce-scatter fusion (ROADMAP item 0): one barrier per step and
    three rotating delta buffers are race free - every rank's global vector is the exact running sum of all
    deltas after every step, under free-running ranks and random pauses."""
    k, n, steps = 4, 24, 40
    w = FakeWorld(k, n, jitter=2e-4, seed=seed)
    rng = np.random.default_rng(50 + seed)
    deltas = rng.integers(-3, 4, size=(steps, k, n)).astype(float)      # integers: sums are exact in fp64
    snaps = [[None] * steps for _ in range(k)]

    def body(rk):
        for t in range(steps):
            rk.fused_scatter_step(deltas[t, rk.me])
            snaps[rk.me][t] = rk.e.copy()
    w.run(body)
    assert _status(w) == [0] * k
    run = np.cumsum(deltas.sum(1), axis=0)
    for r in range(k):
        for t in range(steps):
            np.testing.assert_array_equal(snaps[r][t], run[t])


def test_design_model_two_buffers_are_not_enough():
    """The same protocol with the clear aimed at the buffer of the NEXT step (what two buffers would force) loses
    updates once ranks run freely - the reason for the third buffer."""
    k, n, steps = 4, 8, 20
    lost = False
    for seed in range(8):
        w = FakeWorld(k, n, jitter=3e-4, seed=seed)
        w.rot = [np.zeros((2, n)) for _ in range(k)]

        def step2(rk, delta):
            rk.epoch += 1
            t = rk.epoch
            cur, nxt = t % 2, (t + 1) % 2
            for owner in range(k):
                lo, hi = w.shard(owner)
                w.pause(rk.me)
                with w.lock:
                    w.rot[owner][cur, lo:hi] += delta[lo:hi]
            rk._signal_all(0, t)
            rk._wait_all(0, t)
            for owner in range(k):
                lo, hi = w.shard(owner)
                w.pause(rk.me)
                rk.e[lo:hi] += w.rot[owner][cur, lo:hi]
            lo, hi = w.shard(rk.me)
            if rk.me == 0:
                import time
                time.sleep(0.002)                # a slow owner: its peers are already accumulating step t+1
            w.rot[rk.me][nxt, lo:hi] = 0.0

        def body(rk):
            for t in range(steps):
                step2(rk, np.ones(n))
        w.run(body)
        if any(not np.array_equal(rk.e, np.full(n, float(k * steps))) for rk in w.ranks):
            lost = True
            break
    assert lost
