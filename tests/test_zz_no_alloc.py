"""The allocation-free mini-batch entry point against the regular one, through the C ABI.

`omldm_linear_minibatch_static` is for callers that cannot allocate device memory at call time
(an ingest loop that cannot afford cudaMalloc's device-wide synchronisation, or a host process
whose HBM was fully handed out at start-up).  It keeps its accumulator in a fixed array inside
the library, so the checks here are: same result as `omldm_linear_minibatch` with caller-owned
scratch and as a plain PyTorch fp32 restatement of the update, twice in a row (the shared array
must be clean for the second call), and a clear refusal beyond the documented size limit.

Collected last on purpose (file name): if it fails under `-x`, every other test has already run.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


class Desc(ctypes.Structure):
    _fields_ = [("w", ctypes.c_void_p), ("stats", ctypes.c_void_p), ("d", ctypes.c_int), ("kind", ctypes.c_int),
                ("variant", ctypes.c_int), ("C", ctypes.c_float), ("eps", ctypes.c_float), ("lam", ctypes.c_float)]


def _lib():
    from omldm_b200.ops import _lib as L
    lib = L.kernels()
    lib.omldm_last_error.restype = ctypes.c_char_p
    return lib


def _batch(d, B, seed):
    g = torch.Generator().manual_seed(seed)
    ld = (d + 2 + 31) // 32 * 32
    Z = torch.zeros(B, ld)
    Z[:, :d] = torch.randn(B, d, generator=g)
    w_true = torch.randn(d, generator=g)
    Z[:, d] = 1.0
    Z[:, d + 1] = torch.sign(Z[:, :d] @ w_true)
    return Z.cuda(), ld


@pytest.mark.parametrize("kind,d,B", [(0, 28, 1000), (1, 434, 4096)])
def test_static_scratch_matches_regular_entry_point(kind, d, B):
    lib = _lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    w_ref = torch.zeros(d + 1, device="cuda")
    w_new = torch.zeros(d + 1, device="cuda")
    st_ref = torch.zeros(4, dtype=torch.float64, device="cuda")
    st_new = torch.zeros(4, dtype=torch.float64, device="cuda")
    scratch = torch.zeros(d + 3, device="cuda")          # the regular entry point wants caller scratch
    from omldm_b200.ops import functional as F
    w_pt = torch.zeros(d + 1)                            # the same update in plain PyTorch (fp32, CPU)
    st_pt = torch.zeros(4, dtype=torch.float64)
    for call in range(2):                                # two calls in a row share the built-in array
        Z, ld = _batch(d, B, seed=100 + call)
        with F.use_impl("torch"):
            F.linear_minibatch(Z.cpu(), d, w_pt, st_pt, torch.zeros(d + 3), kind, 1, 0.5, 0.05, 0.01, True)
        ref = Desc(w_ref.data_ptr(), st_ref.data_ptr(), d, kind, 1, 0.5, 0.05, 0.01)
        new = Desc(w_new.data_ptr(), st_new.data_ptr(), d, kind, 1, 0.5, 0.05, 0.01)
        rc = lib.omldm_linear_minibatch(ctypes.c_void_p(Z.data_ptr()), ld, B, ctypes.byref(ref),
                                        ctypes.c_void_p(scratch.data_ptr()), 1, stream)
        assert rc == 0, lib.omldm_last_error()
        rc = lib.omldm_linear_minibatch_static(ctypes.c_void_p(Z.data_ptr()), ld, B, ctypes.byref(new), 1, stream)
        assert rc == 0, lib.omldm_last_error()
        torch.cuda.synchronize()
        assert torch.allclose(w_new, w_ref, rtol=1e-3, atol=1e-5), f"weights differ after call {call}"
        assert torch.allclose(st_new, st_ref, rtol=1e-4), f"statistics differ after call {call}"
        assert torch.allclose(w_new.cpu(), w_pt, rtol=1e-3, atol=1e-5), f"differs from the PyTorch reference after call {call}"
        assert torch.allclose(st_new.cpu()[:3], st_pt[:3], rtol=1e-4), f"statistics differ from PyTorch after call {call}"
    assert float(w_ref.abs().sum()) > 0                  # the update actually happened


def test_static_scratch_rejects_models_beyond_the_limit():
    lib = _lib()
    d = 4096                                             # OMLDM_STATIC_SCRATCH_MAX_D is 4093
    Z, ld = _batch(d, 8, seed=1)
    w = torch.zeros(d + 1, device="cuda")
    st = torch.zeros(4, dtype=torch.float64, device="cuda")
    desc = Desc(w.data_ptr(), st.data_ptr(), d, 0, 1, 0.5, 0.05, 0.01)
    rc = lib.omldm_linear_minibatch_static(ctypes.c_void_p(Z.data_ptr()), ld, 8, ctypes.byref(desc), 1,
                                           ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    assert b"OMLDM_STATIC_SCRATCH_MAX_D" in lib.omldm_last_error()
    torch.cuda.synchronize()
    assert float(w.abs().sum()) == 0.0                   # nothing was launched
