"""Fused protocol kernels on ONE GPU: k ranks emulated by a single cooperative launch
(B200_PROFILING.md: kernels that wait on each other must not be separate launches on one GPU).
Reference = the same reduction done with plain torch ops."""
import ctypes
import math

import pytest
import torch

from omldm_b200.ops import _lib
from omldm_b200.parallel.comm import PAD_FGM, PAD_PHI, PipelineComm

pytestmark = pytest.mark.gpu
DEV = "cuda"
_i, _f, _u, _sz = ctypes.c_int, ctypes.c_float, ctypes.c_uint32, ctypes.c_size_t


def make(k, n):
    comm = PipelineComm(n, DEV, 0, 1, None, emulate=k)
    e = torch.zeros(k, comm.n, device=DEV)
    return comm, e


def K():
    return _lib.kernels()


def sp():
    return _lib.stream_ptr()


@pytest.mark.parametrize("k", [2, 4, 8])
@pytest.mark.parametrize("n", [32, 1000, 1 << 20])
def test_allreduce_mean(k, n):
    comm, e = make(k, n)
    xs = [torch.randn(comm.n, device=DEV) for _ in range(k)]
    for r in range(k):
        comm.bufs[r].copy_(xs[r])
    ref = torch.stack(xs).mean(0)
    for _ in range(2):   # two epochs back to back: flags are monotonic
        _lib.check(K().omldm_sync_allreduce(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0),
                                            _u(comm.next_epoch()), _i(0), sp()), "allreduce")
    torch.cuda.synchronize()
    for r in range(k):
        assert torch.allclose(comm.bufs[r], ref, rtol=1e-5, atol=1e-6)
        assert torch.allclose(e[r], ref, rtol=1e-5, atol=1e-6)
    assert comm.status() == 0
    comm.close()


@pytest.mark.parametrize("k", [2, 8])
def test_allreduce_sum_delta(k):
    n = 4096
    comm, e = make(k, n)
    g0 = torch.randn(comm.n, device=DEV)
    e[:] = g0
    deltas = [torch.randn(comm.n, device=DEV) for _ in range(k)]
    for r in range(k):
        comm.bufs[r].copy_(deltas[r])
    _lib.check(K().omldm_sync_allreduce(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0), _i(1),
                                        _u(comm.next_epoch()), _i(0), sp()), "allreduce")
    torch.cuda.synchronize()
    ref = g0 + torch.stack(deltas).sum(0)
    for r in range(k):
        assert torch.allclose(e[r], ref, rtol=1e-5, atol=1e-5)
        assert float(comm.bufs[r].abs().max()) == 0.0
    comm.close()


def test_async_and_easgd():
    k, n = 4, 2048
    comm, e = make(k, n)
    w0 = torch.randn(comm.n, device=DEV)
    for r in range(k):
        comm.bufs[r].copy_(w0)
        comm.hubs[r].copy_(w0)
        e[r] = w0
    upd = [torch.randn(comm.n, device=DEV) for _ in range(k)]
    for r in range(k):
        comm.bufs[r].add_(upd[r])
    _lib.check(K().omldm_sync_async(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _i(0), sp()),
               "async")
    torch.cuda.synchronize()
    # hub shard owners hold w0 + mean(updates) once every emulated rank has pushed
    ref = w0 + torch.stack(upd).mean(0)
    for r in range(k):
        lo, hi = comm.shard_bounds(r)
        assert torch.allclose(comm.hubs[r][lo:hi], ref[lo:hi], rtol=1e-5, atol=1e-5)
        assert torch.equal(comm.bufs[r], e[r])
    # EASGD: the centre moves toward the workers, the workers toward the centre
    xs = [comm.bufs[r].clone() for r in range(k)]
    centre = torch.cat([comm.hubs[r][slice(*comm.shard_bounds(r))] for r in range(k)])
    _lib.check(K().omldm_sync_easgd(comm.struct_ptr(), _sz(comm.n), _f(0.25), _i(0), sp()), "easgd")
    torch.cuda.synchronize()
    new_centre = torch.cat([comm.hubs[r][slice(*comm.shard_bounds(r))] for r in range(k)])
    # whatever the interleaving, elastic moves conserve  sum_i x_i + centre  and pull both sides together
    total_before = sum(xs) + centre
    total_after = sum(comm.bufs[r] for r in range(k)) + new_centre
    assert torch.allclose(total_before, total_after, rtol=1e-4, atol=1e-4)
    spread_before = sum(float((x - centre).norm()) for x in xs)
    spread_after = sum(float((comm.bufs[r] - new_centre).norm()) for r in range(k))
    assert spread_after < spread_before
    comm.close()


def test_ssp_clock_wait():
    k, n = 4, 1024
    comm, e = make(k, n)
    for c in range(1, 4):
        _lib.check(K().omldm_sync_ssp(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _u(c), _i(1),
                                      _i(0), sp()), "ssp")
    torch.cuda.synchronize()
    assert comm.status() == 0
    from omldm_b200.parallel.comm import PAD_CLOCK
    for r in range(k):
        assert comm.pads[r][PAD_CLOCK:PAD_CLOCK + k].tolist() == [3] * k
    comm.close()


def test_gm_votes():
    k, n = 4, 512
    comm, e = make(k, n)
    for r in range(k):
        comm.bufs[r].fill_(0.001 * (r + 1))
    before = [b.clone() for b in comm.bufs]
    _lib.check(K().omldm_sync_gm(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(10.0),
                                 _u(comm.next_epoch()), _i(0), sp()), "gm")
    torch.cuda.synchronize()
    for r in range(k):   # nobody violated: nothing moves
        assert torch.equal(comm.bufs[r], before[r])
    comm.bufs[2].fill_(1.0)   # rank 2 leaves the ball
    ref = torch.stack([b.clone() for b in comm.bufs]).mean(0)
    _lib.check(K().omldm_sync_gm(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(10.0),
                                 _u(comm.next_epoch()), _i(0), sp()), "gm")
    torch.cuda.synchronize()
    for r in range(k):
        assert torch.allclose(comm.bufs[r], ref, rtol=1e-5) and torch.allclose(e[r], ref, rtol=1e-5)
    assert comm.status() == 0
    comm.close()


def test_fgm_rounds():
    k, n, T = 4, 256, 1.0
    comm, e = make(k, n)
    state = torch.zeros(8 * k, device=DEV)
    state[1::8] = -1.0
    fl = comm.pads[0].view(torch.float32)
    fl[PAD_FGM + 2] = math.sqrt(T) / 2
    fl[PAD_PHI:PAD_PHI + 16] = -math.sqrt(T)
    rounds_seen = 0
    for step in range(1, 121):
        for r in range(k):   # every worker drifts away from the estimate a little per step
            comm.bufs[r].add_(0.001 * (r + 1))
        _lib.check(K().omldm_sync_fgm(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(T),
                                      _f(0.01), _u(step), _lib.ptr(state), _i(0), sp()), "fgm")
        torch.cuda.synchronize()
        rounds = int(state[4].item())
        if rounds > rounds_seen:
            rounds_seen = rounds
            for r in range(1, k):   # after a round every replica equals the new estimate
                assert torch.allclose(comm.bufs[r], comm.bufs[0], rtol=1e-5, atol=1e-6)
                assert torch.allclose(e[r], comm.bufs[r], rtol=1e-5, atol=1e-6)
        # safe-zone invariant: between rounds the average drift stays inside the ball
        avg = torch.stack([comm.bufs[r] - e[r] for r in range(k)]).mean(0)
        assert float(avg.norm()) <= math.sqrt(T) * 1.5   # a few steps of detection lag are inherent
    assert rounds_seen >= 3
    assert comm.status() == 0
    comm.close()


def test_barrier_and_query_reduce():
    k = 8
    comm, _ = make(k, 16)
    _lib.check(K().omldm_device_barrier(comm.struct_ptr(), _u(comm.next_epoch()), sp()), "barrier")
    vals = torch.arange(8 * k, device=DEV, dtype=torch.float32)
    out = torch.zeros(8 * k, device=DEV)
    _lib.check(K().omldm_query_reduce(comm.struct_ptr(), _lib.ptr(vals), _lib.ptr(out), _u(comm.next_epoch()), sp()),
               "query_reduce")
    torch.cuda.synchronize()
    ref = vals.view(k, 8).sum(0)
    assert torch.allclose(out.view(k, 8), ref.expand(k, 8))
    assert comm.status() == 0
    comm.close()
