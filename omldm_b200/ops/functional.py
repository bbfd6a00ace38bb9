"""Python surface of the sm_100a kernels (csrc/*.cu), one function per kernel.

Every op has two implementations with identical semantics:

* native - the hand-written CUDA kernel through the C ABI (mandatory for CUDA tensors unless the
  torch implementation is selected explicitly);
* torch  - plain PyTorch ops.  It is (a) what runs on CPU tensors so host-side logic is testable
  without a GPU, and (b) with `use_impl("torch")` on CUDA tensors, the PyTorch(+cuBLAS) baseline
  the native path is measured against (BASELINE.md "what the build measures itself against").

Staged record rows ("Z rows", pitch ld, multiple of 32 floats):
    [ x_0..x_{d-1} | 1.0 bias/valid (0 = empty row) | target | zero pad ]
"""
from __future__ import annotations

import contextlib
import ctypes
from typing import Optional, Tuple

import torch

from . import _lib
from .gemm import gemm_nn, gemm_outer, gemm_outer_scatter, gemm_tn  # noqa: F401  (re-exported)

_IMPL = "auto"


@contextlib.contextmanager
def use_impl(name: str):
    """Force the implementation: "native", "torch" or "auto" (native on CUDA tensors)."""
    global _IMPL
    old, _IMPL = _IMPL, name
    try:
        yield
    finally:
        _IMPL = old


def set_impl(name: str) -> None:
    global _IMPL
    _IMPL = name


def native(*tensors: torch.Tensor) -> bool:
    if _IMPL == "torch":
        return False
    if _IMPL == "emu":          # CUDA-on-CPU emulation of the kernel sources (_lib.use_emulator), host tensors
        return True
    on_gpu = any(t is not None and t.is_cuda for t in tensors)
    if _IMPL == "native" and not on_gpu:
        raise _lib.NativeError("native implementation requested for CPU tensors")
    return on_gpu


def pad32(n: int) -> int:
    return (n + 31) // 32 * 32


def _k():
    return _lib.kernels()


_i = ctypes.c_int
_f = ctypes.c_float
_p = _lib.ptr


class LinearDesc(ctypes.Structure):
    _fields_ = [("w", ctypes.c_void_p), ("stats", ctypes.c_void_p), ("d", ctypes.c_int),
                ("kind", ctypes.c_int), ("variant", ctypes.c_int), ("C", ctypes.c_float),
                ("eps", ctypes.c_float), ("lam", ctypes.c_float)]


class CommStruct(ctypes.Structure):
    _fields_ = [("bufs", ctypes.c_void_p), ("hubs", ctypes.c_void_p), ("pads", ctypes.c_void_p),
                ("local", ctypes.c_void_p), ("mc_buf", ctypes.c_void_p), ("rank", ctypes.c_int),
                ("world", ctypes.c_int), ("emulate", ctypes.c_int), ("nhub", ctypes.c_int)]


KIND_PA, KIND_PA_REG, KIND_SVM = 0, 1, 2


# ------------------------------------------------------------------------------ ingest
def holdouts_before(c: int) -> int:
    r = c % 10
    return 2 * (c // 10) + (r - 8 if r > 8 else 0)


def holdout_swap(Z: torch.Tensor, ring: torch.Tensor, d: int, count0: int) -> None:
    """In-place 2-in-10 holdout with evict-then-train (FlinkSpoke.scala:92-107)."""
    B, ld = Z.shape
    cap = ring.shape[0]
    if native(Z, ring):
        _lib.check(_k().omldm_holdout_swap(_p(Z), _i(ld), _i(B), _i(d), _p(ring), _i(cap),
                                           ctypes.c_longlong(count0), _lib.stream_ptr()), "holdout_swap")
        return
    h0, h1 = holdouts_before(count0), holdouts_before(count0 + B)
    if h1 == h0:
        return
    # vectorised shift along each slot's chain (same result as the sequential rule).  The index maps
    # depend only on (count0 mod 10, h0 mod cap, B, cap) once the ring is full, so they are built on
    # the host once and cached on the device: the steady state issues no host<->device syncs.
    import numpy as np
    steady = h0 >= cap
    key = (count0 % 10, h0 % cap, B, cap, str(Z.device)) if steady else None
    idx = _HOLDOUT_IDX.get(key) if key is not None else None
    if idx is None:
        h = np.arange(h0, h1)
        rows = (h // 2) * 10 + 8 + (h & 1) - count0             # batch row of holdout h
        prev = h - cap                                           # the holdout it evicts
        in_batch = np.nonzero(prev >= h0)[0]
        from_ring = np.nonzero((prev < h0) & (h >= cap))[0]
        last = np.nonzero(h >= h1 - cap)[0]                      # final occupant of each touched slot
        dev = Z.device
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        idx = (t(rows), t(in_batch), t(prev[in_batch] - h0), t(from_ring), t(h[from_ring] % cap),
               t(h[last] % cap), t(last))
        if key is not None:
            _HOLDOUT_IDX[key] = idx
    rows_t, inb, inb_src, fr, fr_slot, last_slot, last_i = idx
    old_rows = Z[rows_t]
    new_vals = torch.zeros_like(old_rows)
    if inb.numel():
        new_vals[inb] = old_rows[inb_src]
    if fr.numel():
        new_vals[fr] = ring[fr_slot]
    Z[rows_t] = new_vals
    ring[last_slot] = old_rows[last_i]


_HOLDOUT_IDX: dict = {}


def holdout_swap_dev(Z: torch.Tensor, ring: torch.Tensor, d: int, count_dev: torch.Tensor, ordered: bool = True) -> None:
    """Graph-replayable holdout: `count_dev` = int64[2] on the device, [0] the running record count (advanced by the
    kernel itself), [1] the kernel's arrival counter.  `ordered=False`: one exchange per ring slot instead of the shift
    along the chain - same trained set, same ring, for learners that only sum over the batch."""
    B, ld = Z.shape
    cap = ring.shape[0]
    fn = _k().omldm_holdout_swap_dev if ordered else _k().omldm_holdout_exchange_dev
    _lib.check(fn(_p(Z), _i(ld), _i(B), _i(d), _p(ring), _i(cap), _p(count_dev), _lib.stream_ptr()), "holdout_swap_dev")


def curve_sample(stats: torch.Tensor, ring: torch.Tensor, counter: torch.Tensor) -> None:
    """One learning-curve sample (cumulative loss sum, records fitted) into the device ring [cap, 2] (fp64);
    `counter` = int64[1] samples written so far.  Device work only (graph-capturable)."""
    cap = ring.shape[0]
    if native(stats):
        _lib.check(_k().omldm_curve_sample(_p(stats), _p(ring), _i(cap), _p(counter), _lib.stream_ptr()), "curve_sample")
        return
    i = int(counter[0])
    ring[i % cap, 0] = stats[0]
    ring[i % cap, 1] = stats[2]
    counter[0] = i + 1


def stage_rows(raw: torch.Tensor, d: int, ld: Optional[int] = None) -> torch.Tensor:
    """[B, d+1] (features..., target) -> staged rows [B, ld]."""
    B = raw.shape[0]
    ld = ld or pad32(d + 2)
    Z = torch.empty(B, ld, device=raw.device, dtype=torch.float32)
    if native(raw):
        _lib.check(_k().omldm_stage_rows(_p(raw), _i(raw.stride(0)), _i(B), _i(d), _p(Z), _i(ld),
                                         _lib.stream_ptr()), "stage_rows")
        return Z
    Z.zero_()
    Z[:, :d] = raw[:, :d]
    Z[:, d] = 1.0
    Z[:, d + 1] = raw[:, d]
    return Z


def wire_f16_pitch(d: int) -> int:
    """Row pitch (in halves) of the compact fp16 wire format: even(d) features + fp32 target."""
    return ((d + 1) & ~1) + 2


def pack_wire_f16(Z: torch.Tensor, d: int) -> torch.Tensor:
    """Staged fp32 rows (or [B, d+1] raw rows) -> compact wire rows [B, wire_f16_pitch(d)] of fp16.
    Host-side helper for producers and tests; the target keeps its fp32 bits."""
    B = Z.shape[0]
    tgt_col = d + 1 if Z.shape[1] >= d + 2 else d
    out = torch.zeros(B, wire_f16_pitch(d), dtype=torch.float16, device=Z.device)
    out[:, :d] = Z[:, :d].to(torch.float16)
    tgt = torch.empty(B, dtype=torch.float32, device=Z.device).copy_(Z[:, tgt_col])     # unit stride even for B = 1
    out[:, -2:] = tgt.view(torch.float16).reshape(B, 2)
    return out


def stage_rows_f16(raw: torch.Tensor, d: int, ld: Optional[int] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Compact fp16 wire rows -> staged fp32 rows [B, ld] (csrc/ingest.cu::stage_rows_f16_kernel)."""
    B = raw.shape[0]
    ld = ld or pad32(d + 2)
    Z = out if out is not None else torch.empty(B, ld, device=raw.device, dtype=torch.float32)
    if native(raw):
        _lib.check(_k().omldm_stage_rows_f16(_p(raw), _i(raw.stride(0)), _i(B), _i(d), _p(Z), _i(ld),
                                             _lib.stream_ptr()), "stage_rows_f16")
        return Z
    Z.zero_()
    Z[:, :d] = raw[:, :d].to(torch.float32)
    Z[:, d] = 1.0
    Z[:, d + 1] = raw[:, -2:].contiguous().view(torch.float32).reshape(B)
    return Z


# ------------------------------------------------------------------------------ fp16 staged rows ("f16s")
def pad64(n: int) -> int:
    return (n + 63) // 64 * 64


def f16s_pitch(d: int) -> int:
    """Row pitch (in halves) of fp16 staged rows [x | 1 | y_hi | y_lo | 0-pad]: the layout the fp16 Gram kernel
    (csrc/gram_f16.cu) reads in place."""
    return pad64(d + 3)


def pack_f16s(Z: torch.Tensor, d: int) -> torch.Tensor:
    """Staged fp32 rows (or raw [B, d+1] rows) -> fp16 staged rows [B, f16s_pitch(d)].  The target is split into
    two halves (y = y_hi + y_lo keeps 22 significand bits); empty staged rows stay empty."""
    B = Z.shape[0]
    staged = Z.shape[1] >= d + 2
    y = Z[:, d + 1 if staged else d].float()
    out = torch.zeros(B, f16s_pitch(d), dtype=torch.float16, device=Z.device)
    out[:, :d] = Z[:, :d].to(torch.float16)
    out[:, d] = 1.0
    yh = y.to(torch.float16)
    out[:, d + 1] = yh
    out[:, d + 2] = (y - yh.float()).to(torch.float16)
    if staged:
        out[Z[:, d] == 0] = 0
    return out


def expand_rows_f16s(Zh: torch.Tensor, d: int, ld: Optional[int] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp16 staged rows -> fp32 staged rows [B, ld] (csrc/rows_f16.cu::expand_rows_f16s_kernel)."""
    B = Zh.shape[0]
    ld = ld or pad32(d + 2)
    Z = out if out is not None else torch.empty(B, ld, device=Zh.device, dtype=torch.float32)
    if native(Zh):
        _lib.check(_k().omldm_expand_rows_f16s(_p(Zh), _i(Zh.stride(0)), _i(B), _i(d), _p(Z), _i(ld),
                                               _lib.stream_ptr()), "expand_rows_f16s")
        return Z
    Z.zero_()
    Z[:, : d + 1] = Zh[:, : d + 1].float()
    Z[:, d + 1] = Zh[:, d + 1].float() + Zh[:, d + 2].float()
    return Z


def sq_residual_f16s(Zh: torch.Tensor, d: int, wt: torch.Tensor, stats: torch.Tensor, like_gram: bool = False,
                     curve_ring: Optional[torch.Tensor] = None, curve_counter: Optional[torch.Tensor] = None) -> None:
    """stats[0] += sum over valid rows of (w.x + b - y)^2, stats[2] += #valid; wt = [w, b, -1, -1] (d + 3 floats).
    `like_gram`: the pass runs beside the fp16 Gram kernel - walk the rows in the order of its batch-axis splits and
    use two blocks per SM (what fits next to a Gram CTA) so that it stays one wave.
    `curve_ring` [cap, 2] fp64 + `curve_counter` int64[2] (samples written, completion word): the last block also
    writes the learning-curve sample of `curve_sample`."""
    if native(Zh):
        rpr = int(_k().omldm_gram_f16_rows_per_split(_i(Zh.shape[0]), _i(d + 3), _i(0))) \
            if (like_gram and not _lib._emulated) else 0
        cap = curve_ring.shape[0] if curve_ring is not None else 0
        done = ctypes.c_void_p(curve_counter.data_ptr() + 8) if curve_ring is not None else None
        _lib.check(_k().omldm_sq_residual_f16s_tail(_p(Zh), _i(Zh.stride(0)), _i(Zh.shape[0]), _i(d), _p(wt), _p(stats),
                                                    _i(rpr), _i(2 * 148 if like_gram else 0), _p(curve_ring), _i(cap),
                                                    _p(curve_counter) if curve_ring is not None else None, done,
                                                    _lib.stream_ptr()), "sq_residual_f16s")
        return
    valid = Zh[:, d] != 0
    r = Zh[:, : d + 3].float() @ wt[: d + 3]
    stats[0] += (r[valid].double() ** 2).sum()
    stats[2] += valid.sum().double()
    if curve_ring is not None:
        curve_sample(stats, curve_ring, curve_counter)


class GramPeers(ctypes.Structure):
    _fields_ = [("owner_base", ctypes.c_void_p * 16), ("world", ctypes.c_int), ("rank", ctypes.c_int),
                ("rows_per_owner", ctypes.c_int), ("lead", ctypes.c_int)]


class GramLoss(ctypes.Structure):
    """omldm_gram_loss: the prequential loss formed inside the Gram kernel (csrc/gram_f16.cu)."""
    _fields_ = [("wt", ctypes.c_void_p), ("dots", ctypes.c_void_p), ("words", ctypes.c_void_p), ("stats", ctypes.c_void_p),
                ("curve_ring", ctypes.c_void_p), ("curve_counter", ctypes.c_void_p), ("curve_cap", ctypes.c_int),
                ("d", ctypes.c_int)]


def gram_loss_workspace(max_batch: int, device):
    """(dots, words) for `gram_f16(..., loss=)`: zero before the first call, left zero by every call."""
    return (torch.zeros(max_batch + 64, dtype=torch.float32, device=device),
            torch.zeros(3 * 128, dtype=torch.int32, device=device))


def make_gram_loss(wt: torch.Tensor, d: int, stats: torch.Tensor, dots: torch.Tensor, words: torch.Tensor,
                   curve_ring: Optional[torch.Tensor] = None, curve_counter: Optional[torch.Tensor] = None) -> GramLoss:
    g = GramLoss()
    g.wt, g.dots, g.words, g.stats = wt.data_ptr(), dots.data_ptr(), words.data_ptr(), stats.data_ptr()
    g.curve_ring = curve_ring.data_ptr() if curve_ring is not None else None
    g.curve_counter = curve_counter.data_ptr() if curve_ring is not None else None
    g.curve_cap = curve_ring.shape[0] if curve_ring is not None else 0
    g.d = d
    return g


def gram_f16_scratch(device) -> torch.Tensor:
    return torch.zeros(512, dtype=torch.int32, device=device)


def gram_f16(Zh: torch.Tensor, n: int, C: torch.Tensor, scratch: Optional[torch.Tensor] = None,
             peers: Optional[GramPeers] = None, ksplit: int = 0, loss: Optional[GramLoss] = None) -> None:
    """C[:n, :n] (upper 32 x 32 blocks) += Zh[:, :n]^T Zh[:, :n] on the tcgen05 CTA-pair kernel (csrc/gram_f16.cu).
    With `peers`, the rows of C (this worker's running sum) are also copied into the worker's slot at the GPU
    that owns them.  With `loss` (native path only), the same kernel adds the batch's squared residuals and valid-row
    count to the statistics the descriptor names - what `sq_residual_f16s` does in a pass of its own."""
    if native(Zh, C):
        _lib.check(_k().omldm_gram_f16_loss(_p(Zh), _i(Zh.stride(0)), _i(Zh.shape[0]), _i(n), _p(C), _i(C.stride(0)),
                                            ctypes.byref(peers) if peers is not None else None, _p(scratch), _i(ksplit),
                                            ctypes.byref(loss) if loss is not None else None,
                                            _lib.stream_ptr()), "gram_f16")
        return
    if loss is not None:
        raise ValueError("gram_f16: the in-kernel loss exists on the native path only")
    if Zh.is_cuda:      # PyTorch baseline on the GPU: cuBLAS fp16 GEMM with fp32 output, whole matrix
        X = Zh[:, :pad32(n)] if pad32(n) <= Zh.shape[1] else Zh
        G = torch.mm(X.t(), X, out_dtype=torch.float32)
        C[:n, :n] += G[:n, :n]
        return
    X = Zh[:, :n].float()
    G = X.t() @ X
    # the kernel writes the 32 x 32 blocks that touch the upper triangle
    i = torch.arange(n, device=G.device)
    keep = (i[None, :] // 32 * 32 + 31) >= (i[:, None] // 32 * 32)
    C[:n, :n] += G * keep


def densify(indptr: torch.Tensor, indices: torch.Tensor, values: torch.Tensor, targets: Optional[torch.Tensor],
            d: int, ld: Optional[int] = None, out: Optional[torch.Tensor] = None,
            sq_out: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """CSR batch -> staged dense rows and squared row norms (bias included)."""
    B = indptr.numel() - 1
    ld = ld or pad32(d + 2)
    Z = out if out is not None else torch.empty(B, ld, device=values.device, dtype=torch.float32)
    sq = sq_out if sq_out is not None else torch.empty(B, device=values.device, dtype=torch.float32)
    if native(values):
        _lib.check(_k().omldm_densify(_p(indptr), _p(indices), _p(values), _i(B), _i(d), _p(targets), _p(Z),
                                      _i(ld), _p(sq), _lib.stream_ptr()), "densify")
        return Z, sq
    Z.zero_()
    rows = torch.repeat_interleave(torch.arange(B, device=values.device), (indptr[1:] - indptr[:-1]).long())
    ok = (indices >= 0) & (indices < d)
    Z.index_put_((rows[ok], indices[ok].long()), values[ok], accumulate=True)
    sq.zero_()
    sq.index_add_(0, rows[ok], values[ok] ** 2)
    sq += 1.0
    Z[:, d] = 1.0
    if targets is not None:
        Z[:, d + 1] = targets
    return Z, sq


# ------------------------------------------------------------------------------ preprocessors
def poly_features(Zin: torch.Tensor, d_in: int, factors: torch.Tensor, d_out: int) -> torch.Tensor:
    B = Zin.shape[0]
    degree = factors.shape[1]
    ldout = pad32(d_out + 2)
    Zout = torch.empty(B, ldout, device=Zin.device, dtype=torch.float32)
    if native(Zin):
        _lib.check(_k().omldm_poly_features(_p(Zin), _i(Zin.stride(0)), _i(d_in), _p(factors), _i(degree),
                                            _i(d_out), _p(Zout), _i(ldout), _i(B), _lib.stream_ptr()),
                   "poly_features")
        return Zout
    xe = torch.cat([Zin[:, :d_in], torch.ones(B, 1, device=Zin.device)], 1)
    idx = factors.long().clone()
    idx[idx < 0] = d_in
    feats = xe[:, idx.reshape(-1)].reshape(B, d_out, degree).prod(-1)
    valid = Zin[:, d_in:d_in + 1]
    Zout.zero_()
    Zout[:, :d_out] = feats * (valid != 0)
    Zout[:, d_out] = Zin[:, d_in]
    Zout[:, d_out + 1] = Zin[:, d_in + 1]
    return Zout


def standard_scaler(Z: torch.Tensor, d: int, state: torch.Tensor, scratch: torch.Tensor, fit: bool = True) -> None:
    """state: float64 [1 + 2d] = [count, mean, M2]; scratch: float64 [3 + 4d] zeros."""
    B = Z.shape[0]
    if native(Z):
        _lib.check(_k().omldm_standard_scaler(_p(Z), _i(Z.stride(0)), _i(d), _i(B), _p(state), _p(scratch),
                                              _i(1 if fit else 0), _lib.stream_ptr()), "standard_scaler")
        return
    valid = Z[:, d] != 0
    X = Z[valid, :d].double()
    n, mean, m2 = state[0].clone(), state[1:1 + d].clone(), state[1 + d:1 + 2 * d].clone()
    if fit and X.shape[0] > 0:
        nb = float(X.shape[0])
        mb = X.mean(0)
        m2b = ((X - mb) ** 2).sum(0)
        tot = n + nb
        delta = mb - mean
        m2 = m2 + m2b + delta ** 2 * n * nb / tot
        mean = mean + delta * nb / tot
        n = tot
        state[0], state[1:1 + d], state[1 + d:1 + 2 * d] = n, mean, m2
    var = m2 / (n - 1.0) if float(n) > 1 else torch.zeros_like(m2)
    inv = torch.where(var > 1e-24, 1.0 / torch.sqrt(var.clamp_min(1e-300)), torch.ones_like(var))
    Z[valid, :d] = ((Z[valid, :d] - mean.float()) * inv.float())


def minmax_scaler(Z: torch.Tensor, d: int, state: torch.Tensor, fit: bool = True) -> None:
    """state: float32 [2d] = [min, max] (initialise to +inf / -inf)."""
    B = Z.shape[0]
    if native(Z):
        _lib.check(_k().omldm_minmax_scaler(_p(Z), _i(Z.stride(0)), _i(d), _i(B), _p(state), _i(1 if fit else 0),
                                            _lib.stream_ptr()), "minmax_scaler")
        return
    valid = Z[:, d] != 0
    X = Z[valid, :d]
    if fit and X.shape[0] > 0:
        state[:d] = torch.minimum(state[:d], X.min(0).values)
        state[d:] = torch.maximum(state[d:], X.max(0).values)
    rng = state[d:] - state[:d]
    inv = torch.where(rng > 1e-30, 1.0 / rng.clamp_min(1e-30), torch.zeros_like(rng))
    Z[valid, :d] = (X - state[:d]) * inv


# ------------------------------------------------------------------------------ linear learners
def _linear_coef(kind, variant, C, eps, f, y, n2):
    if kind == KIND_PA_REG:
        err = f - y
        loss = (err.abs() - eps).clamp_min(0)
        aux = err * err
        sign = torch.where(err > 0, -1.0, 1.0)
    else:
        yy = torch.where(y > 0, 1.0, -1.0)
        loss = (1 - yy * f).clamp_min(0)
        aux = (yy * f <= 0).float()
        sign = yy
    if variant == 0:
        tau = loss / n2
    elif variant == 1:
        tau = torch.minimum(torch.full_like(loss, C), loss / n2)
    else:
        tau = loss / (n2 + 1.0 / (2.0 * C))
    tau = torch.where((loss > 0) & (n2 > 0), tau, torch.zeros_like(tau))
    return sign * tau, loss, aux


def linear_exact(Z: torch.Tensor, d: int, w: torch.Tensor, stats: torch.Tensor, kind: int, variant: int,
                 C: float, eps: float, lam: float, desc_dev: Optional[torch.Tensor] = None,
                 n_pipelines: int = 1) -> None:
    """Record-by-record update of one (or, with desc_dev, many) linear pipelines.

    w: float32 [d+1]; stats: float64 [4] = [loss, aux, fitted, svm_step].
    """
    B, ld = Z.shape
    if native(Z):
        if desc_dev is None:
            desc = LinearDesc(w.data_ptr(), stats.data_ptr(), d, kind, variant, C, eps, lam)
            desc_dev = torch.frombuffer(bytearray(bytes(desc)), dtype=torch.uint8).to(Z.device)
        _lib.check(_k().omldm_linear_exact(_p(Z), _i(ld), _i(B), _p(desc_dev), _i(n_pipelines), _i(d),
                                           _lib.stream_ptr()), "linear_exact")
        return
    wz = w[:d + 1].double().clone()
    loss_sum = aux_sum = 0.0
    fitted = 0
    t = int(stats[3].item())
    Zc = Z[:, :d + 2].double()
    for r in range(B):
        if Zc[r, d] == 0:
            continue
        x, y = Zc[r, :d + 1], Zc[r, d + 1]
        f = torch.dot(wz, x)
        n2 = torch.dot(x, x)
        if kind == KIND_SVM:
            yy = 1.0 if y > 0 else -1.0
            loss = max(0.0, 1.0 - yy * float(f))
            aux = 1.0 if yy * float(f) <= 0 else 0.0
            t += 1
            eta = 1.0 / (lam * t)
            wz = (1 - eta * lam) * wz + ((eta * yy) * x if yy * float(f) < 1 else 0)
        else:
            c, l, a = _linear_coef(kind, variant, C, eps, f, y, n2)
            loss, aux = float(l), float(a)
            wz = wz + c * x
        loss_sum += loss
        aux_sum += aux
        fitted += 1
    w[:d + 1] = wz.float()
    stats[0] += loss_sum
    stats[1] += aux_sum
    stats[2] += fitted
    stats[3] = t


def linear_exact_multi(Z: torch.Tensor, max_d: int, desc_dev: torch.Tensor, n_pipelines: int) -> None:
    """One launch advances `n_pipelines` linear pipelines (one warp each) over the same batch - the
    multi-tenant case of FlinkSpoke.scala:97-105 where every pipeline sees every record."""
    B, ld = Z.shape
    _lib.check(_k().omldm_linear_exact(_p(Z), _i(ld), _i(B), _p(desc_dev), _i(n_pipelines), _i(max_d),
                                       _lib.stream_ptr()), "linear_exact(multi)")


def make_linear_descs(entries, device) -> torch.Tensor:
    """Pack (w, stats, d, kind, variant, C, eps, lam) tuples into a device descriptor array."""
    raw = bytearray()
    for (w, stats, d, kind, variant, C, eps, lam) in entries:
        raw += bytes(LinearDesc(w.data_ptr(), stats.data_ptr(), d, kind, variant, C, eps, lam))
    return torch.frombuffer(raw, dtype=torch.uint8).to(device)


def linear_minibatch_scratch(d: int, device) -> torch.Tensor:
    return torch.zeros(d + 3, device=device, dtype=torch.float32)


def linear_minibatch(Z: torch.Tensor, d: int, w: torch.Tensor, stats: torch.Tensor, scratch: torch.Tensor,
                     kind: int, variant: int, C: float, eps: float, lam: float, mean_update: bool = True) -> None:
    B, ld = Z.shape
    if native(Z):
        desc = LinearDesc(w.data_ptr(), stats.data_ptr(), d, kind, variant, C, eps, lam)
        _lib.check(_k().omldm_linear_minibatch(_p(Z), _i(ld), _i(B), ctypes.byref(desc), _p(scratch),
                                               _i(1 if mean_update else 0), _lib.stream_ptr()), "linear_minibatch")
        return
    valid = Z[:, d] != 0
    X = Z[valid, :d + 1]
    y = Z[valid, d + 1]
    if X.shape[0] == 0:
        return
    f = X @ w[:d + 1]
    n2 = (X * X).sum(1)
    if kind == KIND_SVM:
        yy = torch.where(y > 0, 1.0, -1.0)
        loss = (1 - yy * f).clamp_min(0)
        aux = (yy * f <= 0).float()
        t = float(stats[3].item()) + 1.0
        eta = 1.0 / (lam * t)
        viol = yy * f < 1
        w[:d + 1] = (1 - eta * lam) * w[:d + 1] + (eta / max(1, B)) * (X[viol].t() @ yy[viol])
        stats[3] += 1
    else:
        c, loss, aux = _linear_coef(kind, variant, C, eps, f, y, n2)
        act = int((c != 0).sum().item())
        if act:
            delta = X.t() @ c
            w[:d + 1] += delta / act if mean_update else delta
    stats[0] += loss.double().sum()
    stats[1] += aux.double().sum()
    stats[2] += X.shape[0]


def linear_predict(Z: torch.Tensor, d: int, w: torch.Tensor) -> torch.Tensor:
    B = Z.shape[0]
    out = torch.empty(B, device=Z.device, dtype=torch.float32)
    if native(Z):
        _lib.check(_k().omldm_linear_predict(_p(Z), _i(Z.stride(0)), _i(B), _p(w), _i(d), _p(out),
                                             _lib.stream_ptr()), "linear_predict")
        return out
    return Z[:, :d] @ w[:d] + w[d]


# ------------------------------------------------------------------------------ ORR helpers
def sq_residual(Z: torch.Tensor, d: int, wt: torch.Tensor, stats: torch.Tensor) -> None:
    """stats[0] += sum over valid rows of (Z_r . wt)^2; stats[2] += #valid.  wt = [w, b, -1]."""
    B = Z.shape[0]
    n = wt.numel()
    if native(Z):
        _lib.check(_k().omldm_sq_residual(_p(Z), _i(Z.stride(0)), _i(B), _i(d), _p(wt), _i(n), _p(stats),
                                          _lib.stream_ptr()), "sq_residual")
        return
    valid = (Z[:, d] != 0)
    r = (Z[:, :n] @ wt) * valid                 # no boolean indexing: stays asynchronous on the GPU
    stats[0] += (r.double() ** 2).sum()
    stats[2] += valid.sum()


def split_tf32(Z: torch.Tensor):
    """z = hi + lo with hi = z truncated to tf32 (what the tensor core keeps of an fp32 operand) and lo the exact
    remainder: hi^T hi + hi^T lo + lo^T hi recovers the fp32 product to ~2^-21 (error-compensated Gram, `precision:
    "tf32x3"`)."""
    hi = (Z.view(torch.int32) & -8192).view(torch.float32)
    return hi, Z - hi


def mirror_upper(A: torch.Tensor, n: int) -> None:
    if native(A):
        _lib.check(_k().omldm_mirror_upper(_p(A), _i(A.stride(0)), _i(n), _lib.stream_ptr()), "mirror_upper")
        return
    iu = torch.triu(A[:n, :n], 1)
    A[:n, :n] = torch.triu(A[:n, :n]) + iu.t()


# ------------------------------------------------------------------------------ sparse rows (config 5)
def sparse_row_pitch(nnz: int) -> int:
    return pad32(2 * nnz + 2)


def sparse_rows(idx: torch.Tensor, val: torch.Tensor, tgt: torch.Tensor, batch: int, nnz: int,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """ELL slot (idx[B*nnz] int32, val[B*nnz], target[B]) -> sparse rows [B, pad32(2 nnz + 2)]:
    [index bits | values | valid = 1 | target | 0-pad].  With the pseudo feature count 2*nnz these obey the
    staged-row conventions (valid flag in column d, target in d + 1), so the holdout kernel applies as is."""
    ld = sparse_row_pitch(nnz)
    rows = out if out is not None else torch.empty(batch, ld, dtype=torch.float32, device=idx.device)
    if native(idx):
        _lib.check(_k().omldm_sparse_rows(_p(idx), _p(val), _p(tgt), _i(batch), _i(nnz), _p(rows), _i(ld),
                                          _lib.stream_ptr()), "sparse_rows")
        return rows
    rows.zero_()
    rows[:, :nnz] = idx.view(batch, nnz).view(torch.float32)
    rows[:, nnz:2 * nnz] = val.view(batch, nnz)
    rows[:, 2 * nnz] = 1.0
    rows[:, 2 * nnz + 1] = tgt
    return rows


def sparse_rows_to_dense(rows: torch.Tensor, nnz: int, d: int) -> torch.Tensor:
    """Sparse rows -> staged dense rows [B, pad32(d + 2)] (duplicate indices accumulate, out-of-range ones are
    dropped, empty rows stay empty).  Used for holdout scoring and as the reference of the sparse kernels."""
    B = rows.shape[0]
    Z = torch.zeros(B, pad32(d + 2), dtype=torch.float32, device=rows.device)
    if B == 0:
        return Z
    col = rows[:, :nnz].contiguous().view(torch.int32).long()
    v = rows[:, nnz:2 * nnz]
    ok = (col >= 0) & (col < d)
    Z.scatter_add_(1, torch.where(ok, col, torch.zeros_like(col)), torch.where(ok, v, torch.zeros_like(v)))
    valid = rows[:, 2 * nnz] != 0
    Z[:, d] = valid.float()
    Z[:, d + 1] = rows[:, 2 * nnz + 1]
    Z[~valid] = 0.0
    return Z


def mcpa_sparse_step(rows: torch.Tensor, nnz: int, W: torch.Tensor, d: int, C: int, variant: int, Creg: float,
                     coef: torch.Tensor, stats: torch.Tensor, active_before: torch.Tensor, mean_update: bool,
                     wt: Optional[torch.Tensor] = None) -> None:
    """One mini-batch MultiClassPA step from sparse rows against the dense model W[C, ldw]
    (csrc/sparse_mcpa.cu).  CUDA only; callers on CPU densify and use the dense step.  `wt`: (d + 1) * 16 floats of
    scratch for the class-minor copy of W the scores are gathered from (C <= 16)."""
    B = rows.shape[0]
    _lib.check(_k().omldm_mcpa_sparse_step_wt(_p(rows), _i(rows.stride(0)), _i(B), _i(nnz), _p(W), _i(W.stride(0)),
                                              _i(d), _i(C), _i(variant), _f(Creg), _p(coef), _p(stats),
                                              _p(active_before), _i(1 if mean_update else 0), _p(wt),
                                              _lib.stream_ptr()), "mcpa_sparse_step")


def ridge_cg_workspace(n: int, device, max_iter: int = 32):
    """(work floats, scalar block, info[2]) for `ridge_cg`; allocate once, reuse for every refresh."""
    work = torch.empty(5 * n, dtype=torch.float32, device=device)
    scal = torch.zeros(4 * (max_iter + 2), dtype=torch.float64, device=device)
    info = torch.zeros(2, dtype=torch.float32, device=device)
    return work, scal, info


def ridge_cg(W: torch.Tensor, n: int, lam: float, x: torch.Tensor, work: torch.Tensor, scal: torch.Tensor,
             info: torch.Tensor, max_iter: int = 32, tol: float = 1e-5, blocks: int = 32) -> None:
    """Solve (A + lam I) x = b in place in `x` (its content is the starting point), where
    A = W[:n, :n] (symmetric, full storage) and b = W[:n, n]: warm-started Jacobi-preconditioned
    conjugate gradients, one cooperative kernel (csrc/ridge_cg.cu).  info <- (iterations, ||r|| / ||b||)."""
    if native(W):
        _lib.check(_k().omldm_ridge_cg(_p(W), _i(W.stride(0)), _i(n), _f(lam), ctypes.c_void_p(W.data_ptr() + 4 * n),
                                       _i(W.stride(0)), _p(x), _p(work), _p(scal), _i(max_iter), _f(tol), _p(info),
                                       _i(blocks), _lib.stream_ptr()), "ridge_cg")
        return
    A = W[:n, :n]
    b = W[:n, n]
    minv = 1.0 / torch.where(A.diagonal() + lam > 0, A.diagonal() + lam, torch.ones_like(A.diagonal()))
    r = b - (A @ x[:n] + lam * x[:n])
    z = minv * r
    p = z.clone()
    rz, rr = float(r.double() @ z.double()), float(r.double() @ r.double())
    bb = float(b.double() @ b.double()) or 1.0
    it = 0
    while it < max_iter and rr > tol * tol * bb:
        q = A @ p + lam * p
        pq = float(q.double() @ p.double())
        if not pq > 0.0:
            break
        alpha = rz / pq
        x[:n] += alpha * p
        r = r - alpha * q
        z = minv * r
        rz_new, rr = float(r.double() @ z.double()), float(r.double() @ r.double())
        p = z + (rz_new / rz if rz > 0 else 0.0) * p
        rz = rz_new
        it += 1
    info[0], info[1] = float(it), (rr / bb) ** 0.5


def ridge_chol(W: torch.Tensor, n: int, lam: float, x: torch.Tensor, info: Optional[torch.Tensor] = None,
               blocks: int = 0) -> torch.Tensor:
    """Solve (A + lam I) x = b directly: A = W[:n, :n] (symmetric, full storage), b = W[n, :n] (row n).  Blocked
    Cholesky in one cooperative kernel (csrc/ridge_chol.cu); W is destroyed.  info[0] <- 1 if a pivot was not positive."""
    if info is None:
        info = torch.zeros(2, dtype=torch.float32, device=W.device)
    if native(W):
        _lib.check(_k().omldm_ridge_chol(_p(W), _i(W.stride(0)), _i(n), _f(lam), _p(x), _p(info), _i(blocks),
                                         _lib.stream_ptr()), "ridge_chol")
        return info
    A = W[:n, :n].double().clone()
    A.diagonal().add_(lam)
    L, bad = torch.linalg.cholesky_ex(A)
    if int(bad) != 0:
        info[0] = 1.0
        x[:n] = float("nan")
        return info
    x[:n] = torch.cholesky_solve(W[n, :n].double()[:, None], L)[:, 0].float()
    info[0] = 0.0
    return info


# ------------------------------------------------------------------------------ k-means
def kmeans_assign(Z, d, S, cnorm, K, T, stats, assign=None) -> None:
    B = Z.shape[0]
    if native(Z):
        _lib.check(_k().omldm_kmeans_assign(_p(Z), _i(Z.stride(0)), _i(d), _i(B), _p(S), _i(S.stride(0)), _p(cnorm),
                                            _i(K), _p(T), _i(T.stride(0)), _p(assign), _p(stats), _lib.stream_ptr()),
                   "kmeans_assign")
        return
    valid = Z[:, d] != 0
    v = cnorm[None, :K] - 2 * S[:, :K]
    best, bj = v.min(1)
    nrm = (Z[:, :d] ** 2).sum(1)
    T.zero_()
    rows = torch.nonzero(valid).squeeze(1)
    T[rows, bj[rows]] = 1.0
    stats[0] += (nrm + best).clamp_min(0)[valid].double().sum()
    stats[2] += int(valid.sum().item())
    if assign is not None:
        assign.copy_(torch.where(valid, bj, torch.full_like(bj, -1)).int())


def kmeans_assign_fused(Z, d, centroids, cnorm, K, T, stats, assign=None) -> bool:
    """Distance GEMM + argmin + one-hot rows in one tcgen05 kernel (csrc/umma_gemm.cu::umma_kmeans_kernel).  Returns
    False when the fused kernel does not apply (CPU tensors, K > 256): use gemm_tn + kmeans_assign then."""
    if not (native(Z) and Z.is_cuda) or K > 256 or _IMPL == "emu":
        return False
    _lib.check(_k().omldm_kmeans_assign_fused(_p(Z), _i(Z.stride(0)), _i(d), _i(Z.shape[0]), _p(centroids),
                                              _i(centroids.stride(0)), _p(cnorm), _i(K), _p(T), _i(T.stride(0)),
                                              _p(assign), _p(stats), _lib.stream_ptr()), "kmeans_assign_fused")
    return True


def kmeans_finalize(sums, sums2, K, d, centroids, cnorm) -> None:
    if native(sums):
        _lib.check(_k().omldm_kmeans_finalize(_p(sums), _p(sums2), _i(sums.stride(0)), _i(K), _i(d), _p(centroids),
                                              _i(centroids.stride(0)), _p(cnorm), _lib.stream_ptr()), "kmeans_finalize")
        return
    tot = sums[:K] + (sums2[:K] if sums2 is not None else 0)
    cnt = tot[:, d]
    inv = torch.where(cnt > 0, 1.0 / cnt.clamp_min(1e-30), torch.zeros_like(cnt))
    centroids.zero_()
    centroids[:K, :d] = tot[:, :d] * inv[:, None]
    cn = (centroids[:K, :d] ** 2).sum(1)
    cnorm[:K] = torch.where(cnt > 0, cn, torch.full_like(cn, float("inf")))


def kmeans_exact(Z, d, sums, K, stats) -> None:
    B = Z.shape[0]
    if native(Z):
        _lib.check(_k().omldm_kmeans_exact(_p(Z), _i(Z.stride(0)), _i(d), _i(B), _p(sums), _i(sums.stride(0)), _i(K),
                                           _p(stats), _lib.stream_ptr()), "kmeans_exact")
        return
    for r in range(B):
        if Z[r, d] == 0:
            continue
        x = Z[r, :d]
        cnt = sums[:K, d]
        empty = torch.nonzero(cnt <= 0)
        if empty.numel():
            j, loss = int(empty[0]), 0.0
        else:
            dist = ((sums[:K, :d] / cnt[:, None] - x) ** 2).sum(1)
            j = int(dist.argmin())
            loss = float(dist[j])
        sums[j, :d] += x
        sums[j, d] += 1
        stats[0] += loss
        stats[2] += 1


# ------------------------------------------------------------------------------ multiclass PA
def row_sqnorm(Z: torch.Tensor, n: int) -> torch.Tensor:
    B = Z.shape[0]
    out = torch.empty(B, device=Z.device, dtype=torch.float32)
    if native(Z):
        _lib.check(_k().omldm_row_sqnorm(_p(Z), _i(Z.stride(0)), _i(n), _i(B), _p(out), _lib.stream_ptr()), "row_sqnorm")
        return out
    return (Z[:, :n] ** 2).sum(1)


def mcpa_coef(Z, d, S, C, sqnorm, variant, Creg, T, stats) -> None:
    """stats: float64 [4] += [loss, mistakes, fitted, active]."""
    B = Z.shape[0]
    if native(Z):
        _lib.check(_k().omldm_mcpa_coef(_p(Z), _i(Z.stride(0)), _i(d), _i(B), _p(S), _i(S.stride(0)), _i(C), _p(sqnorm),
                                        _i(variant), _f(Creg), _p(T), _i(T.stride(0)), _p(stats), _lib.stream_ptr()),
                   "mcpa_coef")
        return
    valid = Z[:, d] != 0
    y = (Z[:, d + 1] + 0.5).long().clamp(0, C - 1)
    Sc = S[:, :C].clone()
    sy = Sc.gather(1, y[:, None]).squeeze(1)
    Sc.scatter_(1, y[:, None], float("-inf"))
    best, rj = Sc.max(1)
    loss = (1 - (sy - best)).clamp_min(0)
    n2 = 2 * sqnorm
    if variant == 0:
        tau = loss / n2
    elif variant == 1:
        tau = torch.minimum(torch.full_like(loss, Creg), loss / n2)
    else:
        tau = loss / (n2 + 1.0 / (2.0 * Creg))
    tau = torch.where((loss > 0) & (n2 > 0) & valid, tau, torch.zeros_like(tau))
    T.zero_()
    ar = torch.arange(B, device=Z.device)
    T[ar, y] = tau
    T[ar, rj] = -tau
    stats[0] += loss[valid].double().sum()
    stats[1] += (best >= sy)[valid].double().sum()
    stats[2] += int(valid.sum().item())
    stats[3] += int((tau > 0).sum().item())


def scale_rows(T: torch.Tensor, C: int, scale_dev: torch.Tensor, reciprocal: bool = False) -> None:
    B = T.shape[0]
    if native(T):
        _lib.check(_k().omldm_scale_rows(_p(T), _i(T.stride(0)), _i(B), _i(C), _p(scale_dev), _i(1 if reciprocal else 0),
                                         _lib.stream_ptr()), "scale_rows")
        return
    s = 1.0 / scale_dev.clamp_min(1.0) if reciprocal else scale_dev
    T[:, :C] *= s


def axpy(y: torch.Tensor, x: torch.Tensor, alpha: float) -> None:
    if native(y):
        _lib.check(_k().omldm_axpy(_p(y), _p(x), _f(alpha), ctypes.c_size_t(y.numel()), _lib.stream_ptr()), "axpy")
        return
    y.add_(x, alpha=alpha)


# ------------------------------------------------------------------------------ neural network glue (csrc/nn.cu)
def nn_hidden_act(H: torch.Tensor, od: int) -> None:
    """In place: relu on the first `od` columns, constant 1 in column `od` (bias input of the next layer)."""
    if native(H):
        _lib.check(_k().omldm_nn_hidden_act(_p(H), _i(H.stride(0)), _i(H.shape[0]), _i(od), _lib.stream_ptr()),
                   "nn_hidden_act")
        return
    H[:, :od].clamp_(min=0)
    H[:, od] = 1.0


def nn_out_grad(out: torch.Tensor, Z: torch.Tensor, d: int, od: int, classes: int, G: torch.Tensor,
                stats: torch.Tensor, n_valid: torch.Tensor) -> None:
    """Loss statistics and un-normalised gradient rows of the output layer (see include/omldm_kernels.h)."""
    B = Z.shape[0]
    if native(out):
        _lib.check(_k().omldm_nn_out_grad(_p(out), _i(out.stride(0)), _p(Z), _i(Z.stride(0)), _i(d), _i(B), _i(od),
                                          _i(classes), _p(G), _i(G.stride(0)), _p(stats), _p(n_valid),
                                          _lib.stream_ptr()), "nn_out_grad")
        return
    valid = Z[:, d] != 0
    y = Z[:, d + 1]
    G.zero_()
    o = out[:, :od]
    if classes > 1:
        p = torch.softmax(o, 1)
        yi = torch.round(y).long().clamp(0, od - 1)
        ll = -torch.log(p.gather(1, yi[:, None]).squeeze(1).clamp_min(1e-30))
        g = p.clone()
        g[torch.arange(B, device=Z.device), yi] -= 1.0
        G[:, :od] = g * valid[:, None]
        stats[0] += (ll * valid).sum().double()
    else:
        err = o[:, 0] - y
        G[:, 0] = 2.0 * err * valid
        stats[0] += (err * err * valid).sum().double()
    stats[2] += valid.sum().double()
    n_valid += valid.sum().float()


def nn_relu_mask(g: torch.Tensor, act: torch.Tensor, n: int) -> None:
    if native(g):
        _lib.check(_k().omldm_nn_relu_mask(_p(g), _i(g.stride(0)), _p(act), _i(act.stride(0)), _i(g.shape[0]), _i(n),
                                           _lib.stream_ptr()), "nn_relu_mask")
        return
    g[:, :n].mul_((act[:, :n] > 0).float())


def nn_sgd_step(W: torch.Tensor, gW: torch.Tensor, lr: float, n_valid: torch.Tensor) -> None:
    """W -= lr / max(n_valid, 1) * gW with n_valid on the device."""
    if native(W):
        _lib.check(_k().omldm_nn_sgd_step(_p(W), _p(gW), ctypes.c_size_t(W.numel()), _f(lr), _p(n_valid),
                                          _lib.stream_ptr()), "nn_sgd_step")
        return
    W.sub_(gW * (lr / n_valid.clamp_min(1.0)))


# ------------------------------------------------------------------------------ fused pipeline (config 3)
def fused_poly_scaler_linear(Z: torch.Tensor, d_in: int, factors: torch.Tensor, d_out: int, scaler_state: torch.Tensor,
                             scaler_scratch: torch.Tensor, w: torch.Tensor, stats: torch.Tensor,
                             lin_scratch: torch.Tensor, kind: int, variant: int, C: float, eps: float, lam: float,
                             mean_update: bool = True) -> None:
    """PolynomialFeatures -> StandardScaler(fit) -> mini-batch linear update without materialising the
    expanded rows (two passes over the raw batch).  Equivalent to the three separate ops."""
    B = Z.shape[0]
    if native(Z):
        desc = LinearDesc(w.data_ptr(), stats.data_ptr(), d_out, kind, variant, C, eps, lam)
        _lib.check(_k().omldm_fused_poly_scaler_linear(
            _p(Z), _i(Z.stride(0)), _i(d_in), _i(B), _p(factors), _i(factors.shape[1]), _i(d_out), _p(scaler_state),
            _p(scaler_scratch), ctypes.byref(desc), _p(lin_scratch), _i(1 if mean_update else 0), _lib.stream_ptr()),
            "fused_poly_scaler_linear")
        return
    Zp = poly_features(Z, d_in, factors, d_out)
    standard_scaler(Zp, d_out, scaler_state, scaler_scratch, True)
    linear_minibatch(Zp, d_out, w, stats, lin_scratch, kind, variant, C, eps, lam, mean_update)


# ------------------------------------------------------------------------------ Hoeffding tree
def _ht_route_torch(Z, feature, threshold, left, right):
    node = torch.zeros(Z.shape[0], dtype=torch.long, device=Z.device)
    ar = torch.arange(Z.shape[0], device=Z.device)
    for _ in range(64):
        f = feature[node].long()
        inner = f >= 0
        if not bool(inner.any()):
            break
        x = Z[ar, f.clamp_min(0)]
        nxt = torch.where(x <= threshold[node], left[node].long(), right[node].long())
        node = torch.where(inner, nxt, node)
    return node


def ht_class(y: torch.Tensor, C: int) -> torch.Tensor:
    c = torch.round(y).long()
    if C == 2:
        c = torch.where(c < 0, torch.zeros_like(c), c)
    return c.clamp(0, C - 1)


def ht_split(feature, threshold, left, right, n_nodes_dev, C, d, cnt, sums, sq, checked, grace, delta, tie,
             n_points, decision, n_splits) -> None:
    """Hoeffding split test and tree growth on the device (csrc/tree.cu::ht_split_eval_kernel /
    ht_split_apply_kernel): no host round trip.  Native only; the torch path keeps the host oracle's split test."""
    cap = feature.shape[0]
    _lib.check(_k().omldm_ht_split(_p(feature), _p(threshold), _p(left), _p(right), _p(n_nodes_dev), _i(cap), _i(C),
                                   _i(d), _p(cnt), _p(sums), _p(sq), _p(checked), _f(grace), _f(delta), _f(tie),
                                   _i(n_points), _p(decision), _p(n_splits), _lib.stream_ptr()), "ht_split")


def ht_accumulate(Z, d, feature, threshold, left, right, slot, C, cnt, sums, sq, stats, leaf_out=None) -> None:
    """Route every record to its leaf and add it to the leaf's (count, sum, sum-of-squares) statistics."""
    B = Z.shape[0]
    if native(Z):
        _lib.check(_k().omldm_ht_accumulate(_p(Z), _i(Z.stride(0)), _i(d), _i(B), _p(feature), _p(threshold), _p(left),
                                            _p(right), _p(slot), _i(C), _p(cnt), _p(sums), _p(sq), _p(leaf_out),
                                            _p(stats), _lib.stream_ptr()), "ht_accumulate")
        return
    valid = Z[:, d] != 0
    node = _ht_route_torch(Z, feature, threshold, left, right)
    s = slot[node].long()
    c = ht_class(Z[:, d + 1], C)
    best, pred = cnt[s].max(1)
    pred = torch.where(best > 0, pred, torch.full_like(pred, -1))
    stats[0] += ((pred != c) & valid).sum()
    stats[2] += valid.sum()
    idx = (s * C + c)[valid]
    cnt.view(-1).index_add_(0, idx, torch.ones(idx.numel(), device=Z.device))
    X = Z[valid, :d]
    sums.view(-1, d).index_add_(0, idx, X)
    sq.view(-1, d).index_add_(0, idx, X * X)
    if leaf_out is not None:
        leaf_out.copy_(torch.where(valid, node, torch.full_like(node, -1)).int())


def ht_route(Z, feature, threshold, left, right) -> torch.Tensor:
    B = Z.shape[0]
    if native(Z):
        out = torch.empty(B, dtype=torch.int32, device=Z.device)
        _lib.check(_k().omldm_ht_route(_p(Z), _i(Z.stride(0)), _i(B), _p(feature), _p(threshold), _p(left), _p(right),
                                       _p(out), _lib.stream_ptr()), "ht_route")
        return out.long()
    return _ht_route_torch(Z, feature, threshold, left, right)
