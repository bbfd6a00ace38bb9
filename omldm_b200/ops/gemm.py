"""tcgen05/TMEM GEMM ops (csrc/umma_gemm.cu).

`gemm_tn(A, B)`      -> A[M,K] @ B[N,K]^T      (score / distance shape)
`gemm_outer(A, B)`   -> A[K,M]^T @ B[K,N]      (Gram / outer-product-update shape)

fp32 tensors, tf32 tensor-core math, fp32 accumulate.  On CPU tensors a plain torch matmul is
used so the host-side logic stays testable without a GPU.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

ACCUMULATE = 1
UPPER_ONLY = 2
MN_SLAB2D = 4
NO_TMA_EPILOGUE = 16


import os as _os
_ENV_FLAGS = (MN_SLAB2D if _os.environ.get('OMLDM_MN_LOAD', '3d') == '2d' else 0) | \
    (NO_TMA_EPILOGUE if _os.environ.get('OMLDM_GEMM_EPILOGUE', 'tma') == 'red' else 0)


def _check2d(t: torch.Tensor, name: str) -> None:
    if t.dim() != 2 or t.dtype != torch.float32 or t.stride(1) != 1:
        raise ValueError(f"{name} must be a 2-D fp32 tensor with unit inner stride")


def gemm_tn(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor | None = None, *,
            accumulate: bool = False, ksplit: int = 1) -> torch.Tensor:
    """out[M,N] (+)= a[M,K] @ b[N,K]^T."""
    _check2d(a, "a"); _check2d(b, "b")
    M, K = a.shape
    N = b.shape[0]
    if b.shape[1] != K:
        raise ValueError("inner dimensions differ")
    if out is None:
        out = torch.zeros(M, N, device=a.device, dtype=torch.float32) if accumulate else \
            torch.empty(M, N, device=a.device, dtype=torch.float32)
    _check2d(out, "out")
    if not _lib.use_native(a, b, out):
        r = a @ b.t()
        out.add_(r) if accumulate else out.copy_(r)
        return out
    flags = (ACCUMULATE if (accumulate or ksplit > 1) else 0) | _ENV_FLAGS
    if ksplit > 1 and not accumulate:
        out.zero_()
    rc = _lib.kernels().omldm_gemm_tn(
        _lib.ptr(a), ctypes.c_int(a.stride(0)), _lib.ptr(b), ctypes.c_int(b.stride(0)),
        _lib.ptr(out), ctypes.c_int(out.stride(0)), ctypes.c_int(M), ctypes.c_int(N),
        ctypes.c_int(K), ctypes.c_int(ksplit), ctypes.c_int(flags), _lib.stream_ptr())
    _lib.check(rc, "omldm_gemm_tn")
    return out


def gemm_nn(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, *, accumulate: bool = False) -> torch.Tensor:
    """out[M,N] (+)= a[M,K] @ b[K,N], both row-major; b needs a pitch that is a multiple of 32 floats."""
    _check2d(a, "a"); _check2d(b, "b"); _check2d(out, "out")
    M, K = a.shape
    N = out.shape[1]
    if b.shape[0] != K or b.shape[1] < N:
        raise ValueError("inner dimensions differ")
    if not _lib.use_native(a, b, out):
        r = a @ b[:, :N]
        out.add_(r) if accumulate else out.copy_(r)
        return out
    rc = _lib.kernels().omldm_gemm_nn(
        _lib.ptr(a), ctypes.c_int(a.stride(0)), _lib.ptr(b), ctypes.c_int(b.stride(0)),
        _lib.ptr(out), ctypes.c_int(out.stride(0)), ctypes.c_int(M), ctypes.c_int(N),
        ctypes.c_int(K), ctypes.c_int(1), ctypes.c_int((ACCUMULATE if accumulate else 0) | (_ENV_FLAGS & NO_TMA_EPILOGUE)),
        _lib.stream_ptr())
    _lib.check(rc, "omldm_gemm_nn")
    return out


def gemm_outer(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor | None = None, *,
               accumulate: bool = False, ksplit: int = 1, upper_only: bool = False,
               m: int | None = None, n: int | None = None) -> torch.Tensor:
    """out[M,N] (+)= a[K,:M]^T @ b[K,:N].  Row pitches of a and b must be multiples of 32 floats.

    `upper_only` skips output tiles strictly below the diagonal (symmetric Gram updates).
    """
    _check2d(a, "a"); _check2d(b, "b")
    K = a.shape[0]
    if b.shape[0] != K:
        raise ValueError("row counts differ")
    M = a.shape[1] if m is None else m
    N = b.shape[1] if n is None else n
    if out is None:
        out = torch.zeros(M, N, device=a.device, dtype=torch.float32)
        accumulate_eff = True
    else:
        accumulate_eff = accumulate
    _check2d(out, "out")
    if not _lib.use_native(a, b, out):
        # torch / cuBLAS implementation (CPU tests and the GPU baseline): every tile is computed
        if accumulate_eff:
            out[:M, :N].addmm_(a[:, :M].t(), b[:, :N])
        else:
            torch.mm(a[:, :M].t(), b[:, :N], out=out[:M, :N]) if out[:M, :N].is_contiguous() else \
                out[:M, :N].copy_(a[:, :M].t() @ b[:, :N])
        return out
    flags = (ACCUMULATE if (accumulate_eff or ksplit > 1) else 0) | (UPPER_ONLY if upper_only else 0)
    if _ENV_FLAGS:
        flags |= _ENV_FLAGS
    if ksplit > 1 and not accumulate_eff:
        out.zero_()
    rc = _lib.kernels().omldm_gemm_outer(
        _lib.ptr(a), ctypes.c_int(a.stride(0)), _lib.ptr(b), ctypes.c_int(b.stride(0)),
        _lib.ptr(out), ctypes.c_int(out.stride(0)), ctypes.c_int(M), ctypes.c_int(N),
        ctypes.c_int(K), ctypes.c_int(ksplit), ctypes.c_int(flags), _lib.stream_ptr())
    _lib.check(rc, "omldm_gemm_outer")
    return out


def gemm_outer_scatter(a: torch.Tensor, b: torch.Tensor, peer_ptrs, rows_per_owner: int, ldc: int, *,
                       ksplit: int = 1, upper_only: bool = False, m: int | None = None, n: int | None = None) -> None:
    """C (+)= a[K,:M]^T @ b[K,:N] with the rows of C reduced straight into their owners' buffers: `peer_ptrs[p]` is
    the (peer-mapped) base address of GPU p's accumulation matrix with pitch `ldc`, which receives the rows
    [p * rows_per_owner, (p + 1) * rows_per_owner).  CUDA only (csrc/umma_gemm.cu::umma_gemm_scatter_kernel)."""
    _check2d(a, "a"); _check2d(b, "b")
    K = a.shape[0]
    M = a.shape[1] if m is None else m
    N = b.shape[1] if n is None else n
    arr = (ctypes.c_void_p * len(peer_ptrs))(*[int(p) for p in peer_ptrs])
    flags = ACCUMULATE | (UPPER_ONLY if upper_only else 0)
    rc = _lib.kernels().omldm_gemm_outer_scatter(
        _lib.ptr(a), ctypes.c_int(a.stride(0)), _lib.ptr(b), ctypes.c_int(b.stride(0)), arr,
        ctypes.c_int(len(peer_ptrs)), ctypes.c_int(rows_per_owner), ctypes.c_int(ldc), ctypes.c_int(M),
        ctypes.c_int(N), ctypes.c_int(K), ctypes.c_int(ksplit), ctypes.c_int(flags), _lib.stream_ptr())
    _lib.check(rc, "omldm_gemm_outer_scatter")
