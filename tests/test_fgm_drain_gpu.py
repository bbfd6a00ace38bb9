"""FGM end-of-stream drain handshake on the device (k ranks emulated in one cooperative launch); the protocol
itself is modelled and stress-tested on CPU in tests/test_fake_backend.py."""
import ctypes
import math

import pytest
import torch

from omldm_b200.ops import _lib
from omldm_b200.parallel.comm import PAD_FGM, PAD_PHI, PipelineComm

pytestmark = pytest.mark.gpu
_i, _f, _u, _sz = ctypes.c_int, ctypes.c_float, ctypes.c_uint32, ctypes.c_size_t


def test_drain_terminates_and_leaves_no_round_open():
    k, n, T = 4, 256, 1.0
    comm = PipelineComm(n, "cuda", 0, 1, None, emulate=k)
    e = torch.zeros(k, comm.n, device="cuda")
    state = torch.zeros(8 * k, device="cuda")
    state[1::8] = -1.0
    fl = comm.pads[0].view(torch.float32)
    fl[PAD_FGM + 2] = math.sqrt(T) / 2
    fl[PAD_PHI:PAD_PHI + 16] = -math.sqrt(T)
    K, sp = _lib.kernels(), _lib.stream_ptr()
    step = 0
    for gen in (1, 2):                                   # two stream segments, two flush generations
        for _ in range(40):
            step += 1
            for r in range(k):
                comm.bufs[r].add_(0.004 * (r + 1))
            _lib.check(K.omldm_sync_fgm(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(T), _f(0.01),
                                        _u(step), _lib.ptr(state), _i(0), sp), "fgm")
        launches = 0
        while True:
            step += 1
            launches += 1
            _lib.check(K.omldm_sync_fgm_drain(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(T),
                                              _f(0.01), _u(step), _lib.ptr(state), _u(gen), _i(0), sp), "fgm_drain")
            torch.cuda.synchronize()
            term = [int(comm.pads[r][PAD_FGM + 5].item()) for r in range(k)]
            if all(t == gen for t in term):
                break
            assert launches < 8, f"no terminate word after {launches} drain steps: {term}"
        assert comm.status() == 0
        rounds = state.view(k, 8)[:, 0].tolist()
        assert len(set(rounds)) == 1                      # every rank executed the same rounds
        for r in range(k):                                # nothing announced is left open
            assert int(comm.pads[r][PAD_FGM + 3].item()) == int(rounds[0])
    assert state[4].item() >= 1
    comm.close()
