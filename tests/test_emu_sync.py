"""The fused protocol kernels of csrc/sync.cu executed on the CPU (tools/cuda_emu): k ranks emulated inside one
launch whose blocks all run concurrently as host threads, sys-scope loads / stores / atomics mapped to host
atomics with the same ordering.  Same checks as tests/test_sync_gpu.py, plus the FGM drain handshake."""
import ctypes
import math

import pytest
import torch

from omldm_b200.ops import functional as F

from omldm_b200.ops import _lib
from omldm_b200.parallel.comm import PAD_CLOCK, PAD_FGM, PAD_PHI, PipelineComm

_i, _f, _u, _sz = ctypes.c_int, ctypes.c_float, ctypes.c_uint32, ctypes.c_size_t
DEV = "cpu"


@pytest.fixture(scope="module", autouse=True)
def emulator():
    if torch.cuda.is_available():
        pytest.skip("the emulator is for machines without a GPU")
    import os
    _lib.use_emulator(asan=bool(os.environ.get("OMLDM_EMU_ASAN")))
    yield
    _lib.stop_emulator()


def make(k, n):
    comm = PipelineComm(n, DEV, 0, 1, None, emulate=k)
    return comm, torch.zeros(k, comm.n)


def K():
    return _lib.kernels()


def sp():
    return _lib.stream_ptr()


@pytest.mark.parametrize("k", [2, 4])
@pytest.mark.parametrize("n", [32, 1000, 5000])
def test_allreduce_mean(k, n):
    comm, e = make(k, n)
    xs = [torch.randn(comm.n) for _ in range(k)]
    for r in range(k):
        comm.bufs[r].copy_(xs[r])
    ref = torch.stack(xs).mean(0)
    for _ in range(2):   # two epochs back to back: flags are monotonic
        _lib.check(K().omldm_sync_allreduce(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0),
                                            _u(comm.next_epoch()), _i(2), sp()), "allreduce")
    for r in range(k):
        assert torch.allclose(comm.bufs[r], ref, rtol=1e-5, atol=1e-6)
        assert torch.allclose(e[r], ref, rtol=1e-5, atol=1e-6)
    assert comm.status() == 0


def test_allreduce_sum_delta_device_managed_epochs():
    k, n = 3, 4096
    comm, e = make(k, n)
    g0 = torch.randn(comm.n)
    e[:] = g0
    tot = torch.zeros(comm.n)
    for step in range(3):                       # epoch argument 0: the kernel keeps its own sequence numbers
        deltas = [torch.randn(comm.n) for _ in range(k)]
        for r in range(k):
            comm.bufs[r].copy_(deltas[r])
        tot += torch.stack(deltas).sum(0)
        _lib.check(K().omldm_sync_allreduce(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0), _i(1), _u(0), _i(2),
                                            sp()), "allreduce")
        for r in range(k):
            assert torch.allclose(e[r], g0 + tot, rtol=1e-5, atol=1e-5)
            assert float(comm.bufs[r].abs().max()) == 0.0
    assert comm.status() == 0


def test_async_and_easgd():
    k, n = 4, 2048
    comm, e = make(k, n)
    w0 = torch.randn(comm.n)
    for r in range(k):
        comm.bufs[r].copy_(w0)
        comm.hubs[r].copy_(w0)
        e[r] = w0
    upd = [torch.randn(comm.n) for _ in range(k)]
    for r in range(k):
        comm.bufs[r].add_(upd[r])
    _lib.check(K().omldm_sync_async(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _i(1), sp()), "async")
    ref = w0 + torch.stack(upd).mean(0)
    for r in range(k):
        lo, hi = comm.shard_bounds(r)
        assert torch.allclose(comm.hubs[r][lo:hi], ref[lo:hi], rtol=1e-5, atol=1e-5)
        assert torch.equal(comm.bufs[r], e[r])
    xs = [comm.bufs[r].clone() for r in range(k)]
    centre = torch.cat([comm.hubs[r][slice(*comm.shard_bounds(r))] for r in range(k)])
    _lib.check(K().omldm_sync_easgd(comm.struct_ptr(), _sz(comm.n), _f(0.25), _i(1), sp()), "easgd")
    new_centre = torch.cat([comm.hubs[r][slice(*comm.shard_bounds(r))] for r in range(k)])
    assert torch.allclose(sum(xs) + centre, sum(comm.bufs[r] for r in range(k)) + new_centre, rtol=1e-4, atol=1e-4)
    assert sum(float((comm.bufs[r] - new_centre).norm()) for r in range(k)) < sum(float((x - centre).norm()) for x in xs)


def test_ssp_clock_wait():
    k, n = 4, 1024
    comm, e = make(k, n)
    for c in range(1, 4):
        _lib.check(K().omldm_sync_ssp(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _u(c), _i(1),
                                      _i(1), sp()), "ssp")
    assert comm.status() == 0
    for r in range(k):
        assert comm.pads[r][PAD_CLOCK:PAD_CLOCK + k].tolist() == [3] * k


def test_gm_votes():
    k, n = 4, 512
    comm, e = make(k, n)
    for r in range(k):
        comm.bufs[r].fill_(0.001 * (r + 1))
    before = [b.clone() for b in comm.bufs]
    _lib.check(K().omldm_sync_gm(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(10.0),
                                 _u(comm.next_epoch()), _i(1), sp()), "gm")
    for r in range(k):   # nobody violated: nothing moves
        assert torch.equal(comm.bufs[r], before[r])
    comm.bufs[2].fill_(1.0)
    ref = torch.stack([b.clone() for b in comm.bufs]).mean(0)
    _lib.check(K().omldm_sync_gm(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(10.0),
                                 _u(comm.next_epoch()), _i(1), sp()), "gm")
    for r in range(k):
        assert torch.allclose(comm.bufs[r], ref, rtol=1e-5) and torch.allclose(e[r], ref, rtol=1e-5)
    assert comm.status() == 0


def _fgm_setup(k, n, T):
    comm, e = make(k, n)
    state = torch.zeros(8 * k)
    state[1::8] = -1.0
    fl = comm.pads[0].view(torch.float32)
    fl[PAD_FGM + 2] = math.sqrt(T) / 2
    fl[PAD_PHI:PAD_PHI + 16] = -math.sqrt(T)
    return comm, e, state


@pytest.mark.parametrize("blocks", [1, 2])
def test_fgm_rounds(blocks):
    k, n, T = 4, 256, 1.0
    comm, e, state = _fgm_setup(k, n, T)
    rounds_seen = 0
    for step in range(1, 81):
        for r in range(k):
            comm.bufs[r].add_(0.0015 * (r + 1))
        _lib.check(K().omldm_sync_fgm(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(T),
                                      _f(0.01), _u(step), _lib.ptr(state), _i(blocks), sp()), "fgm")
        rounds = int(state[4].item())
        if rounds > rounds_seen:
            rounds_seen = rounds
            for r in range(1, k):   # after a round every replica equals the new estimate
                assert torch.allclose(comm.bufs[r], comm.bufs[0], rtol=1e-5, atol=1e-6)
                assert torch.allclose(e[r], comm.bufs[r], rtol=1e-5, atol=1e-6)
        assert len(set(state.view(k, 8)[:, 0].tolist())) == 1, "ranks executed different numbers of rounds"
        avg = torch.stack([comm.bufs[r] - e[r] for r in range(k)]).mean(0)
        assert float(avg.norm()) <= math.sqrt(T) * 1.5
    assert rounds_seen >= 3
    assert comm.status() == 0


def test_fgm_drain_terminates_and_leaves_no_round_open():
    k, n, T = 4, 256, 1.0
    comm, e, state = _fgm_setup(k, n, T)
    step = 0
    for gen in (1, 2):
        for _ in range(30):
            step += 1
            for r in range(k):
                comm.bufs[r].add_(0.004 * (r + 1))
            _lib.check(K().omldm_sync_fgm(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(T), _f(0.01),
                                          _u(step), _lib.ptr(state), _i(1), sp()), "fgm")
        launches = 0
        while True:
            step += 1
            launches += 1
            _lib.check(K().omldm_sync_fgm_drain(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(T),
                                                _f(0.01), _u(step), _lib.ptr(state), _u(gen), _i(1), sp()), "fgm_drain")
            term = [int(comm.pads[r][PAD_FGM + 5].item()) for r in range(k)]
            if all(t == gen for t in term):
                break
            assert launches < 8, f"no terminate word after {launches} drain steps: {term}"
        assert comm.status() == 0
        rounds = state.view(k, 8)[:, 0].tolist()
        assert len(set(rounds)) == 1
        for r in range(k):
            assert int(comm.pads[r][PAD_FGM + 3].item()) == int(rounds[0])
    assert state[4].item() >= 1


@pytest.mark.parametrize("queries_per_step", [0.25, 3])
def test_fgm_rounds_interleaved_with_queries(queries_per_step):
    """FGM round-end rendezvous and the query / termination barrier advance independent counters (round number vs
    query number).  They used to share the ENTRY / EXIT flag words, so after two rounds a query fell through its
    barrier (or, with more queries than rounds, a round-end wait did).  Both orders: fewer queries than rounds and
    more queries than rounds; every query must return the exact sum and no timeout bit may be set."""
    k, n, T = 4, 256, 1.0
    comm, e, state = _fgm_setup(k, n, T)
    vals = torch.zeros(8 * k)
    out = torch.zeros(8 * k)
    nq = 0
    acc = 0.0
    for step in range(1, 61):
        for r in range(k):
            comm.bufs[r].add_(0.004 * (r + 1))
        _lib.check(K().omldm_sync_fgm(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(T),
                                      _f(0.01), _u(0), _lib.ptr(state), _i(1), sp()), "fgm")
        acc += queries_per_step
        while acc >= 1.0:
            acc -= 1.0
            nq += 1
            vals.copy_(torch.arange(8 * k, dtype=torch.float32) + nq)
            out.zero_()
            _lib.check(K().omldm_query_reduce(comm.struct_ptr(), _lib.ptr(vals), _lib.ptr(out), _u(0), sp()), "query")
            assert torch.equal(out.view(k, 8), vals.view(k, 8).sum(0).expand(k, 8)), f"query {nq} at step {step}"
            if nq % 5 == 0:
                _lib.check(K().omldm_device_barrier(comm.struct_ptr(), _u(0), sp()), "barrier")
        assert comm.status() == 0, f"timeout bit at step {step}"
        assert len(set(state.view(k, 8)[:, 0].tolist())) == 1
    assert state[4].item() >= 3 and nq >= 10


def test_barrier_and_query_reduce():
    k = 8
    comm, _ = make(k, 16)
    _lib.check(K().omldm_device_barrier(comm.struct_ptr(), _u(comm.next_epoch()), sp()), "barrier")
    vals = torch.arange(8 * k, dtype=torch.float32)
    out = torch.zeros(8 * k)
    _lib.check(K().omldm_query_reduce(comm.struct_ptr(), _lib.ptr(vals), _lib.ptr(out), _u(comm.next_epoch()), sp()),
               "query_reduce")
    assert torch.allclose(out.view(k, 8), vals.view(k, 8).sum(0).expand(k, 8))
    assert comm.status() == 0


@pytest.mark.parametrize("k,blocks", [(2, 1), (4, 1), (3, 2)])
def test_scatter_push_gather_three_rotating_copies(k, blocks):
    """Reduce-scatter moved into the producer (sync_scatter_push / sync_scatter_gather): after every step each rank's
    global vector is the exact running sum of all deltas; device-managed step and epoch counters; a barrier-type
    kernel in between (a query) must not disturb the rotation."""
    n = 96 * k                                   # model size; the symmetric buffers hold three copies
    comm = PipelineComm(3 * n, DEV, 0, 1, None, emulate=k)
    e = torch.zeros(k, n)
    delta = torch.zeros(k, n)
    run = torch.zeros(n)
    g = torch.Generator().manual_seed(k)
    for step in range(7):
        d = torch.randint(-3, 4, (k, n), generator=g).float()       # integers: every partial sum is exact
        d[:, 5:9] = 0.0                                             # zero vectors are skipped by the push
        delta.copy_(d)
        run += d.sum(0)
        _lib.check(K().omldm_sync_scatter_push(comm.struct_ptr(), _lib.ptr(delta), _sz(n), _u(0), _i(blocks), sp()), "push")
        _lib.check(K().omldm_sync_scatter_gather(comm.struct_ptr(), _lib.ptr(e), _sz(n), _f(1.0), _u(0), _u(0),
                                                 _i(blocks), sp()), "gather")
        for r in range(k):
            assert torch.equal(e[r], run), f"step {step}, rank {r}"
        assert float(delta.abs().max()) == 0.0                      # the push clears the local delta
        if step == 3:                                               # consumes an epoch, not a rotation step
            _lib.check(K().omldm_device_barrier(comm.struct_ptr(), _u(0), sp()), "barrier")
    assert comm.status() == 0
    # at rest exactly one copy per owner still holds data (the one the next-but-one step clears)
    per = (n // 4 + k - 1) // k * 4
    for r in range(k):
        copies = comm.bufs[r].view(3, n)[:, r * per:(r + 1) * per]
        assert int((copies.abs().sum(1) > 0).sum()) <= 1


@pytest.mark.parametrize("k,h", [(4, 1), (4, 2), (3, 2)])
def test_hub_parallelism_fewer_shards_than_workers(k, h):
    """HubParallelism = h < k (FlinkSpoke.scala:181-190): only ranks 0 .. h-1 own a hub shard.  The all-reduce still
    yields the exact mean on every rank, and the asynchronous push lands in the h owners' shards only."""
    n = 1000
    comm = PipelineComm(n, DEV, 0, 1, None, emulate=k, n_hubs=h)
    assert comm.n_hubs == h and comm.shard_bounds(h - 1)[1] == comm.n and comm.shard_bounds(k - 1)[0] == comm.n or h == k
    e = torch.zeros(k, comm.n)
    xs = [torch.randn(comm.n) for _ in range(k)]
    for r in range(k):
        comm.bufs[r].copy_(xs[r])
    _lib.check(K().omldm_sync_allreduce(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _u(0), _i(2), sp()),
               "allreduce")
    ref = torch.stack(xs).mean(0)
    for r in range(k):
        assert torch.allclose(comm.bufs[r], ref, rtol=1e-5, atol=1e-6)
    # asynchronous pushes: hub shards live on the first h ranks only
    for r in range(k):
        comm.hubs[r].zero_()
        comm.bufs[r].fill_(float(r + 1))
    e.zero_()
    _lib.check(K().omldm_sync_async(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0), _i(0), _i(2), sp()), "async")
    tot = float(sum(range(1, k + 1)))
    for r in range(k):
        lo, hi = comm.shard_bounds(r)
        if r < h:
            assert torch.allclose(comm.hubs[r][lo:hi], torch.full((hi - lo,), tot))
        assert float(comm.hubs[r][:lo].abs().max() if lo else 0.0) == 0.0
        assert float(comm.hubs[r][hi:].abs().max() if hi < comm.n else 0.0) == 0.0
    assert comm.status() == 0


@pytest.mark.parametrize("k,blocks,n", [(2, 1, 61), (4, 2, 100), (3, 1, 128)])
def test_sharded_hub_push_and_gather(k, blocks, n):
    """Sharded-hub protocol of the ORR Gram statistics (sync_gram_push_kernel / sync_gram_gather_kernel, driver class
    parallel/fused.py::ShardedGramDriver) with k emulated ranks: every rank copies its running sum into its own slot at
    the owners without waiting; the pull at a refresh sums the slots at the owners and leaves the same global matrix
    (upper 32 x 32 blocks) on every rank.  Several pushes between two pulls (newer values replace older ones) and a
    query barrier in between must not disturb it."""
    from omldm_b200.parallel.fused import ShardedGramDriver
    ld = F.pad64(n)
    nb, nh = ShardedGramDriver.region_floats(ld, k)
    comm = PipelineComm(nb, DEV, 0, 1, None, emulate=k, hub_floats=nh)
    drv = ShardedGramDriver(comm, comm.buf, "sum", {"syncBlocks": blocks}, impl="fused", ld=ld, n=n)
    assert drv.rows_per_owner % 32 == 0 and drv.rows_per_owner * k >= n
    acc = drv.acc.view(k, ld, ld)
    total = torch.zeros(ld, ld)
    g = torch.Generator().manual_seed(k + n)
    i = torch.arange(ld)
    keep = ((i[None, :] // 32 * 32 + 31) >= (i[:, None] // 32 * 32)) & (i[:, None] < n) & (i[None, :] < n)
    for rnd in range(3):
        for step in range(1 + rnd):
            dlt = torch.randint(-3, 4, (k, ld, ld), generator=g).float() * keep        # integers: exact sums
            acc += dlt
            total += dlt.sum(0)
            drv.push()
        drv.gather()
        for r in range(k):
            got = comm.bufs[r].view(ld, ld)
            assert torch.equal(got * keep, total), f"round {rnd}, rank {r}"
        if rnd == 1:
            _lib.check(K().omldm_device_barrier(comm.struct_ptr(), _u(0), sp()), "barrier")
    assert comm.status() == 0


@pytest.mark.parametrize("k,blocks", [(2, 1), (4, 2)])
def test_single_learner_forwarding(k, blocks):
    """sync_forward_kernel: every emulated worker stores its batch into rank 0's buffer (slot seq & 1, position = rank);
    rank 0 sees all k batches of the step, the other slot still holds the previous step."""
    n = 8 * 32                                   # floats per batch (8 staged rows of pitch 32)
    comm = PipelineComm(2 * k * n, DEV, 0, 1, None, emulate=k)
    prev = None
    for seq in range(1, 6):
        Z = torch.randn(k, n)
        _lib.check(K().omldm_sync_forward(comm.struct_ptr(), _lib.ptr(Z), _sz(n), _u(seq), _i(blocks), sp()), "forward")
        slots = comm.bufs[0].view(2, k, n)
        assert torch.equal(slots[seq & 1], Z)
        if prev is not None:
            assert torch.equal(slots[(seq & 1) ^ 1], prev)
        prev = Z
        for r in range(k):                      # rank 0 announced what it has consumed
            from omldm_b200.parallel.comm import PAD_EXIT
            assert int(comm.pads[r][PAD_EXIT]) == seq - 1
    assert comm.status() == 0


def test_heartbeat_and_stall_detection():
    """heartbeat_kernel publishes every emulated rank's counter in every pad; the monitor flags a peer whose counter
    stops advancing (fake clock)."""
    from omldm_b200.parallel.comm import PAD_HEART
    from omldm_b200.utils.watchdog import HeartbeatMonitor
    k = 4
    comm, _ = make(k, 16)
    for beat in (1, 2, 7):
        comm.heartbeat(beat)
        for r in range(k):
            assert comm.pads[r][PAD_HEART:PAD_HEART + k].tolist() == [beat] * k
    assert comm.heartbeats() == [7] * k
    now = [0.0]
    mon = HeartbeatMonitor(k, rank=0, stall_after_s=5.0, clock=lambda: now[0])
    beats = [0, 0, 0, 0]
    for t in range(1, 13):
        now[0] = float(t)
        for r in (0, 1, 3):                       # rank 2 stops beating after t = 4
            beats[r] = t
        if t <= 4:
            beats[2] = t
        mon.observe(list(beats))
        assert mon.stalled() == ([2] if t > 9 else [])


def test_fgm_kernel_and_python_model_take_the_same_decisions():
    """The Python model of the protocol (parallel/fake_backend.py, used for the interleaving stress tests) and the CUDA
    kernel (emulated) are driven with the same drift sequence in lock-step.  The hub polls phi values that the other
    ranks publish concurrently, so a decision may legitimately fall one step earlier or later in either system; beyond
    that the two must agree: same number of rounds (+-1), round ends within two steps of each other, replicas equal
    after every round - so what the model proves carries over to the kernel."""
    import numpy as np
    from omldm_b200.parallel.fake_backend import FakeWorld
    k, n, T = 4, 64, 1.0
    comm, e, state = _fgm_setup(k, n, T)
    w = FakeWorld(k, comm.n)
    w.ranks[0].fgm_init_hub(T)
    rng = np.random.default_rng(5)
    kernel_rounds, model_rounds = [], []
    for step in range(1, 201):
        drift = rng.uniform(0.0005, 0.004, size=(k, 1)) * np.ones((k, comm.n))
        for r in range(k):
            comm.bufs[r].add_(torch.from_numpy(drift[r]).float())
            w.bufs[r][:] += drift[r]
        _lib.check(K().omldm_sync_fgm(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(T),
                                      _f(0.01), _u(step), _lib.ptr(state), _i(1), sp()), "fgm")
        done = []
        w.run(lambda rk: done.append(rk.fgm_step(T)))
        assert len(set(done)) == 1
        if done[0]:
            model_rounds.append(step)
            for r in range(1, k):
                assert np.allclose(w.bufs[r], w.bufs[0])
        if int(state[4].item()) > len(kernel_rounds):
            kernel_rounds.append(step)
            for r in range(1, k):
                assert torch.allclose(comm.bufs[r], comm.bufs[0], rtol=1e-5, atol=1e-6)
    assert len(kernel_rounds) >= 2 and abs(len(kernel_rounds) - len(model_rounds)) <= 1, (kernel_rounds, model_rounds)
    for a, b in zip(kernel_rounds, model_rounds):
        assert abs(a - b) <= 2 * (1 + kernel_rounds.index(a)), (kernel_rounds, model_rounds)
    assert comm.status() == 0


@pytest.mark.parametrize("k", [2, 4])
def test_fgm_driver_runs_the_multigpu_check_scenario(k):
    """scripts/multigpu_check.py's FGM case through the real driver class (parallel/fused.py::FGMDriver): device-managed
    step counter (step argument 0), 150 steps of per-rank drift, drain flush; then a second stream on the same driver."""
    from omldm_b200.parallel.fused import make_driver
    comm = PipelineComm(256, DEV, 0, 1, None, emulate=k)
    e = torch.zeros(k, comm.n)
    drv = make_driver("FGM", comm, e, "mean", {"threshold": 1.0}, "fused")
    for gen in (1, 2):
        for s in range(150):
            for r in range(k):
                comm.bufs[r].add_(0.001 * (r + 1))
            drv.step()
        drv.flush()
        assert drv.flush_timeouts == 0 and comm.status() == 0
        assert drv.device_rounds() >= 2 * gen
        rounds = drv.state.view(k, 8)[:, 0].tolist()
        assert len(set(rounds)) == 1, rounds
        for r in range(k):
            assert int(comm.pads[r][PAD_FGM + 5].item()) == gen
            assert int(comm.pads[r][PAD_FGM + 3].item()) == int(rounds[0])      # no round left open


def test_driver_classes_on_emulated_kernels():
    """The Python driver classes (parallel/fused.py) as the engine uses them on hardware - device-managed epochs /
    clocks (sequence argument 0), default block counts - against the checks of scripts/multigpu_check.py."""
    from omldm_b200.parallel.fused import ScatterSynchronousDriver, make_driver
    k, n = 3, 1000
    # Synchronous, averaged model: two steps back to back
    comm, e = make(k, n)
    xs = [torch.randn(comm.n) for _ in range(k)]
    for r in range(k):
        comm.bufs[r].copy_(xs[r])
    drv = make_driver("Synchronous", comm, e, "mean", {}, "fused")
    assert drv.impl == "fused"
    drv.step(); drv.step()
    ref = torch.stack(xs).mean(0)
    for r in range(k):
        assert torch.allclose(comm.bufs[r], ref, rtol=1e-5, atol=1e-6) and torch.allclose(e[r], ref, rtol=1e-5, atol=1e-6)
    assert comm.status() == 0
    # Synchronous, additive statistics: e <- e + sum of deltas, deltas cleared
    comm, e = make(k, n)
    g0 = torch.randn(comm.n)
    e[:] = g0
    ds = [torch.randn(comm.n) for _ in range(k)]
    for r in range(k):
        comm.bufs[r].copy_(ds[r])
    drv = make_driver("Synchronous", comm, e, "sum", {}, "fused")
    drv.step()
    for r in range(k):
        assert torch.allclose(e[r], g0 + sum(ds), rtol=1e-5, atol=1e-4) and float(comm.bufs[r].abs().max()) == 0
    # Asynchronous: the hub shards hold the mean of all pushes, every worker installed a hub value
    comm, e = make(k, n)
    drv = make_driver("Asynchronous", comm, e, "mean", {}, "fused")
    for r in range(k):
        comm.bufs[r].add_(float(r + 1))
    drv.step()
    for r in range(k):
        lo, hi = comm.shard_bounds(r)
        assert torch.allclose(comm.hubs[r][lo:hi], torch.full((hi - lo,), sum(range(1, k + 1)) / k))
        assert torch.equal(comm.bufs[r], e[r])
    # SSP: clocks advance without a timeout
    comm, e = make(k, n)
    drv = make_driver("SSP", comm, e, "mean", {"staleness": 1}, "fused")
    for _ in range(5):
        for r in range(k):
            comm.bufs[r].add_(0.1)
        drv.step()
    assert comm.status() == 0
    for r in range(k):
        assert comm.pads[r][PAD_CLOCK:PAD_CLOCK + k].tolist() == [5] * k
    # EASGD: what the workers lose the centre gains
    comm, e = make(k, n)
    drv = make_driver("EASGD", comm, None, "mean", {"alpha": 0.25}, "fused")
    for r in range(k):
        comm.bufs[r].fill_(float(r + 1))
    total = sum(comm.bufs[r].clone() for r in range(k))
    drv.step()
    centre = torch.cat([comm.hubs[r][slice(*comm.shard_bounds(r))] for r in range(k)])
    assert torch.allclose(sum(comm.bufs[r] for r in range(k)) + centre, total, rtol=1e-5, atol=1e-5)
    assert float(centre.abs().min()) > 0
    # GM: nothing moves inside the ball, a violation triggers the averaging round
    comm, e = make(k, n)
    drv = make_driver("GM", comm, e, "mean", {"threshold": 10.0}, "fused")
    for r in range(k):
        comm.bufs[r].fill_(0.001 * (r + 1))
    before = [b.clone() for b in comm.bufs]
    drv.step()
    assert all(torch.equal(comm.bufs[r], before[r]) for r in range(k))
    comm.bufs[k - 1].fill_(1.0)
    ref = torch.stack([b.clone() for b in comm.bufs]).mean(0)
    drv.step()
    assert all(torch.allclose(comm.bufs[r], ref, rtol=1e-5) for r in range(k))
    assert comm.status() == 0
    # scatter-fused Synchronous: private deltas, three rotating copies
    m = 96 * k
    comm = PipelineComm(3 * m, DEV, 0, 1, None, emulate=k)
    e, delta, run = torch.zeros(k, m), torch.zeros(k, m), torch.zeros(m)
    drv = make_driver("Synchronous", comm, e, "sum", {"syncFusion": "scatter"}, "fused", scatter_delta=delta)
    assert isinstance(drv, ScatterSynchronousDriver) and not drv.gemm_fused
    g = torch.Generator().manual_seed(1)
    for _ in range(5):
        d = torch.randint(-3, 4, (k, m), generator=g).float()
        delta.copy_(d)
        run += d.sum(0)
        drv.step()
        assert all(torch.equal(e[r], run) for r in range(k)) and float(delta.abs().max()) == 0.0
    assert comm.status() == 0
