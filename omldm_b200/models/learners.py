"""Device learners: the eight learners of PipelineMap.scala:68 on staged record rows.

Every learner keeps its synchronisable state in ONE flat fp32 vector that lives in the
symmetric arena (`PipelineComm.buf`), so the protocol kernels can move it without packing:

  merge "mean": buf = the model itself (PA / RegressorPA / SVM / MultiClassPA / NN), e = last
                global estimate.
  merge "sum" : buf = additive delta since the last sync (ORR Gram matrix, k-means sums/counts),
                e = global statistics; derived quantities (ORR weights, centroids) are refreshed
                after every sync.

The maths mirrors `omldm_b200/oracle/learners.py` (fp64 oracle); kernels are in csrc/.
"""
from __future__ import annotations

from typing import Any, Dict, Optional

import numpy as np
import torch

from ..ops import functional as F
from ..oracle import learners as O
from ..oracle.learners import _h, parse_variant


def pad4(n: int) -> int:
    return (n + 3) // 4 * 4


class GpuLearner:
    name = ""
    task = "regression"
    merge_mode = "mean"

    def __init__(self, hyper: Optional[dict], device):
        self.hyper = dict(hyper or {})
        self.device = torch.device(device)
        self.d: Optional[int] = None
        self.mode = str(_h(hyper, ["trainingMode", "mode"], "minibatch"))
        self.stats = torch.zeros(4, dtype=torch.float64, device=self.device)  # loss, aux, fitted, extra
        self.buf: Optional[torch.Tensor] = None
        self.e: Optional[torch.Tensor] = None
        self.curve_ring: Optional[torch.Tensor] = None      # [cap, 2] fp64 (cumulative loss sum, fitted) per mini-batch
        self.curve_count: Optional[torch.Tensor] = None
        self._curve_sampled = False

    def enable_curve(self, cap: int = 1024) -> None:
        """Test mode: keep a device ring of per-mini-batch (loss sum, fitted) samples (FlinkHub.scala:101-113)."""
        self.curve_ring = torch.zeros(cap, 2, dtype=torch.float64, device=self.device)
        self.curve_count = torch.zeros(2, dtype=torch.int64, device=self.device)   # [samples written, completion word of a fused sample]

    def sample_curve(self) -> None:
        """One sample on the current stream; learners that run their loss pass on a side stream call it there."""
        if self.curve_ring is not None:
            F.curve_sample(self.stats, self.curve_ring, self.curve_count)
            self._curve_sampled = True

    def learning_curve(self):
        """[(mean loss of the mini-batch, records fitted so far)] in time order, from the device ring."""
        if self.curve_ring is None:
            return []
        n = int(self.curve_count[0].item())
        cap = self.curve_ring.shape[0]
        ring = self.curve_ring.cpu()
        idx = [i % cap for i in range(max(0, n - cap), n)]
        out, prev = [], (0.0, 0.0) if n <= cap else None
        for i in idx:
            L, m = float(ring[i, 0]), float(ring[i, 1])
            if prev is not None and m > prev[1]:
                out.append(((L - prev[0]) / (m - prev[1]), int(m)))
            prev = (L, m)
        return out

    # ---- sizing / binding -------------------------------------------------------------
    def n_flat(self, d: int) -> int:
        raise NotImplementedError

    def needs_e(self) -> bool:
        return True

    def bind(self, d: int, buf: torch.Tensor, e: Optional[torch.Tensor]) -> None:
        """`buf` is the symmetric flat vector (zero-filled), `e` a private vector of equal size."""
        self.d, self.buf, self.e = d, buf, e
        self.ld = F.pad32(d + 2)
        self._init()

    def _init(self) -> None: ...

    # ---- training / inference -----------------------------------------------------------
    def fit_batch(self, Z: torch.Tensor) -> None:
        raise NotImplementedError

    def predict(self, Z: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError

    def after_sync(self) -> None:
        """Refresh derived state once the protocol installed a new global model (device work only)."""

    def host_tick(self) -> None:
        """Host-side cadence work after a step (may synchronise); never part of a CUDA graph."""

    def graph_ready(self) -> bool:
        """True once `fit_batch` enqueues the same device work every step (graph-capturable)."""
        return True

    def score(self, Z: torch.Tensor) -> float:
        """accuracy / RMSE / mean inertia on staged holdout rows (empty rows are ignored)."""
        if Z.shape[0] == 0:
            return 0.0
        valid = Z[:, self.d] != 0
        if not bool(valid.any()):
            return 0.0
        p = self.predict(Z)[valid]
        y = Z[valid, self.d + 1]
        if self.task == "classification":
            return float((self._labels(p) == self._labels_true(y)).float().mean().item())
        return float(torch.sqrt(((p - y) ** 2).mean()).item())

    def _labels(self, p):
        return torch.where(p > 0, 1.0, -1.0)

    def _labels_true(self, y):
        return torch.where(y > 0, 1.0, -1.0)

    # ---- bookkeeping ----------------------------------------------------------------------
    def counters(self) -> Dict[str, float]:
        s = self.stats.tolist()
        return {"loss_sum": s[0], "aux": s[1], "fitted": int(s[2]), "extra": s[3]}

    def to_oracle(self) -> O.OracleLearner:
        raise NotImplementedError

    def from_oracle(self, o: O.OracleLearner) -> None:
        raise NotImplementedError

    def to_parameters(self) -> Dict[str, Any]:
        return self.to_oracle().to_parameters()

    def data_structure(self) -> Optional[Dict[str, Any]]:
        return None

    def state_dict(self) -> dict:
        return {"d": self.d, "buf": None if self.buf is None else self.buf.detach().cpu().clone(),
                "e": None if self.e is None else self.e.detach().cpu().clone(), "stats": self.stats.cpu().clone()}

    def load_state_dict(self, sd: dict) -> None:
        if sd["buf"] is not None:
            self.buf.copy_(sd["buf"].to(self.device))
        if sd["e"] is not None and self.e is not None:
            self.e.copy_(sd["e"].to(self.device))
        self.stats.copy_(sd["stats"].to(self.device))
        self.after_sync()


# ====================================================================================== linear
class GpuLinear(GpuLearner):
    """PA classifier / PA regressor / Pegasos SVM (csrc/linear_learners.cu)."""
    kind = F.KIND_PA

    def __init__(self, hyper, device):
        super().__init__(hyper, device)
        self.C = float(_h(hyper, ["C", "c"], 0.01))
        self.variant = parse_variant(hyper)
        self.eps = float(_h(hyper, ["epsilon", "eps"], 0.1))
        self.lam = float(_h(hyper, ["lambda", "lam", "regularization"], 0.01))
        self.mean_update = str(_h(hyper, ["miniBatchUpdate"], "mean")) == "mean"

    def n_flat(self, d):
        return pad4(d + 1)

    def _init(self):
        self.w = self.buf[: self.d + 1]
        self.scratch = F.linear_minibatch_scratch(self.d, self.device)
        self._desc = None

    def fit_batch(self, Z):
        if self.mode == "exact":
            if self._desc is None and Z.is_cuda:
                self._desc = F.make_linear_descs([(self.w, self.stats, self.d, self.kind, self.variant, self.C,
                                                   self.eps, self.lam)], self.device)
            F.linear_exact(Z, self.d, self.w, self.stats, self.kind, self.variant, self.C, self.eps, self.lam,
                           desc_dev=self._desc if F.native(Z) else None)
        else:
            F.linear_minibatch(Z, self.d, self.w, self.stats, self.scratch, self.kind, self.variant, self.C,
                               self.eps, self.lam, self.mean_update)

    def predict(self, Z):
        return F.linear_predict(Z, self.d, self.w)

    def _oracle_cls(self):
        return O.PAClassifier

    def to_oracle(self):
        o = self._oracle_cls()(self.hyper)
        o.ensure_init(self.d)
        o.w = self.w.detach().double().cpu().numpy().copy()
        c = self.counters()
        o.fitted, o.cum_loss = c["fitted"], c["loss_sum"]
        if hasattr(o, "t"):
            o.t = int(c["extra"])
        return o

    def from_oracle(self, o):
        self.w.copy_(torch.from_numpy(np.asarray(o.w, dtype=np.float32)).to(self.device))


class GpuPA(GpuLinear):
    name, task, kind = "PA", "classification", F.KIND_PA


class GpuRegressorPA(GpuLinear):
    name, task, kind = "RegressorPA", "regression", F.KIND_PA_REG

    def _oracle_cls(self):
        return O.PARegressor


class GpuSVM(GpuLinear):
    name, task, kind = "SVM", "classification", F.KIND_SVM

    def _oracle_cls(self):
        return O.SVM


# ====================================================================================== ORR
class GpuORR(GpuLearner):
    """Online ridge regression on the augmented Gram matrix G = sum z^T z.

    Two data paths, one state.  fp16 staged rows z = [x, 1, y_hi, y_lo] go through the CTA-pair tcgen05 kernel
    (csrc/gram_f16.cu) exactly as they arrive from the host; fp32 staged rows z = [x, 1, y] through the tf32 GEMM
    (csrc/umma_gemm.cu).  G is (d + 3) x (d + 3) with pitch ldg = pad64(d + 3): the y_lo row / column is zero on the
    fp32 path and folded into the y row / column before any solve, so both paths (and a mix) give the same system.
    A - lambda I, b and y^T y are all blocks of G, so replicas merge by addition.

    The weights are refreshed every `solveEvery` mini-batches on a side stream by warm-started conjugate gradients
    (csrc/ridge_cg.cu); a refresh that misses its tolerance, and every query, uses the blocked Cholesky of
    csrc/ridge_chol.cu.  No library solver on this path.
    """
    name, task, merge_mode = "ORR", "regression", "sum"

    def __init__(self, hyper, device):
        super().__init__(hyper, device)
        self.lam = float(_h(hyper, ["lambda", "lam", "regularization"], 1.0))
        self.solve_every = int(_h(hyper, ["solveEvery"], 16))
        self._since = 0
        self.solves = 0
        self._side = None
        self._deltas = None
        self._drain = None
        self._fork = None
        self.overlap = str(_h(hyper, ["overlapLoss"], "true")).lower() in ("1", "true", "yes")
        # fp16 rows on a GPU: form the prequential loss inside the Gram kernel (csrc/gram_f16.cu) instead of a pass
        # of its own beside it
        self.loss_in_gram = str(_h(hyper, ["lossInGram"], "false")).lower() in ("1", "true", "yes")
        self._gl_ws = None
        self._gl_desc = None
        self._pending = None
        self._install_at = 0
        self._w_next = None
        # weight refresh: "cg" (warm-started preconditioned conjugate gradients, csrc/ridge_cg.cu, Cholesky when it does
        # not converge) or "cholesky" (csrc/ridge_chol.cu every time)
        self.solver = str(_h(hyper, ["solver"], "cg")).lower()
        self.cg_iters = int(_h(hyper, ["cgMaxIter"], 32))
        self.cg_tol = float(_h(hyper, ["cgTol"], 1e-5))
        self.cg_failures = 0
        self.cg_last = (0, 0.0)          # (iterations, relative residual) of the last checked CG refresh
        self._cg = None
        self._cg_host = None
        self._cg_evt = None
        self._scatter = None
        self._hub = None                 # ShardedGramDriver when the hub is sharded over the GPUs
        self._gram_scratch = None
        self._chol_info = None
        # "tf32x3": error-compensated Gram on the fp32 path (hi / lo split of the rows, 3 tensor-core passes)
        self.precision = str(_h(hyper, ["precision"], "default")).lower()

    def n_flat(self, d):
        ld = F.pad64(d + 3)
        return ld * ld

    def _init(self):
        self.ldg = ld = F.pad64(self.d + 3)
        self.dG = self.buf.view(ld, ld)
        self.G = self.e.view(ld, ld)
        self.wt = torch.zeros(self.d + 3, device=self.device)   # [w, b, -1, -1]
        self.wt[self.d + 1:] = -1.0
        self._work = torch.empty(ld, ld, device=self.device)
        self._chol_info = torch.zeros(2, device=self.device)
        n = self.d + 2
        tiles = sum(1 for mi in range((n + 127) // 128) for nj in range((n + 255) // 256)
                    if not (128 * mi > 256 * nj + 255))
        self.ksplit = max(1, 148 // max(tiles, 1))
        if self.device.type == "cuda":
            self._gram_scratch = F.gram_f16_scratch(self.device)
            self._side = torch.cuda.Stream(device=self.device)
            self._fork = torch.cuda.Stream(device=self.device)
            self._w_next = torch.zeros(self.d + 1, device=self.device)
            # Run the real refresh path once on both streams now (G = 0, so A = lambda I): lazily loaded kernels,
            # workspaces and stream-private allocator blocks are created here instead of inside somebody's timed
            # region.  Not a collective (the sharded hub is attached after bind()).
            lam0, self.lam = self.lam, max(self.lam, 1.0)     # keep the warm-up system non-singular
            self.solve()
            self.solve_async()
            self._install()
            self.lam = lam0
            self.wt[: self.d + 1].zero_()
            self._since = 0
            self.solves = 0
            self.cg_failures = 0
            torch.cuda.synchronize(self.device)

    # ---- synchronisation variants ---------------------------------------------------------------------
    def use_scatter_gemm(self, driver) -> None:
        """`syncFusion: "gemm"`: the tf32 Gram update reduces straight into the shard owners' rotating copies
        (ops.gemm.gemm_outer_scatter) instead of the local delta."""
        k = driver.k
        if self.ldg % k or (self.ldg // k) % 32:
            raise ValueError(f"syncFusion=gemm needs ld / world to be a multiple of 32 (ld = {self.ldg}, world = {k})")
        self._scatter = driver

    def use_sharded_hub(self, driver) -> None:
        """`syncFusion: "hub"` (parallel/fused.py::ShardedGramDriver): `dG` is this worker's private running sum, whose
        touched rows the Gram kernel itself copies into the worker's slots at the owners; `G` is the worker's copy of
        the global statistics, valid after a `gather()`."""
        self._hub = driver

    def _loss_desc(self, B: int):
        """Descriptor of the in-kernel loss (workspace sized for the largest batch seen; rebuilt when the learning-curve
        ring appears or the batch grows)."""
        key = (B, None if self.curve_ring is None else self.curve_ring.data_ptr(), self.wt.data_ptr())
        if self._gl_desc is None or self._gl_desc[0][0] < B or self._gl_desc[0][1:] != key[1:]:
            if self._gl_ws is None or self._gl_ws[0].numel() < B + 64:
                self._gl_ws = F.gram_loss_workspace(B, self.device)
            self._gl_desc = (key, F.make_gram_loss(self.wt, self.d, self.stats, self._gl_ws[0], self._gl_ws[1],
                                                   self.curve_ring, self.curve_count))
        return self._gl_desc[1]

    def _gram(self, Z, ks, n, with_loss: bool = False):
        if Z.dtype == torch.float16:
            hub = self._hub
            loss = self._loss_desc(Z.shape[0]) if with_loss else None
            if hub is not None and Z.is_cuda:
                F.gram_f16(Z, self.d + 3, self.dG, scratch=hub.scratch, peers=hub.peers, loss=loss)
            else:
                F.gram_f16(Z, self.d + 3, self.dG, scratch=self._gram_scratch, loss=loss)
                if hub is not None:
                    hub.push()
            if with_loss:
                self._curve_sampled = self.curve_ring is not None
            return
        if self._scatter is not None and Z.is_cuda:
            F.gemm_outer_scatter(Z, Z, self._scatter.peer_copy_ptrs(), self.ldg // self._scatter.k, self.ldg,
                                 ksplit=ks, upper_only=True, m=n, n=n)
            return
        if self.precision == "tf32x3" and F.native(Z):
            # error-compensated products: z = hi + lo with hi = tf32(z); hi^T hi + hi^T lo + lo^T hi (upper part,
            # so the third product lands in the transposed position and is mirrored by the solve)
            hi, lo = F.split_tf32(Z)
            F.gemm_outer(hi, hi, self.dG, accumulate=True, ksplit=ks, upper_only=True, m=n, n=n)
            F.gemm_outer(hi, lo, self.dG, accumulate=True, ksplit=ks, upper_only=True, m=n, n=n)
            F.gemm_outer(lo, hi, self.dG, accumulate=True, ksplit=ks, upper_only=True, m=n, n=n)
        else:
            F.gemm_outer(Z, Z, self.dG, accumulate=True, ksplit=ks, upper_only=True, m=n, n=n)
        if self._hub is not None:
            self._hub.push()

    def _loss(self, Z):
        if Z.dtype == torch.float16:     # the learning-curve sample rides in the pass's last block
            F.sq_residual_f16s(Z, self.d, self.wt, self.stats, like_gram=self.overlap and Z.is_cuda,
                               curve_ring=self.curve_ring, curve_counter=self.curve_count)
            self._curve_sampled = self.curve_ring is not None
            return
        F.sq_residual(Z, self.d, self.wt[: self.d + 2], self.stats)   # prequential squared error with current w
        self.sample_curve()          # on the loss pass's stream: beside the Gram kernel, not after it

    def fit_batch(self, Z):
        n = self.d + 2
        ks = max(1, min(self.ksplit, Z.shape[0] // 64))
        if Z.is_cuda and Z.dtype == torch.float16 and self.loss_in_gram and F.native(Z) and not F._lib._emulated:
            self._gram(Z, ks, n, with_loss=True)     # one kernel: Gram update, hub push and prequential loss
            return
        if Z.is_cuda and self.overlap:
            # fork: the prequential-loss pass (HBM-bound) runs beside the Gram GEMM (tensor-bound);
            # they touch disjoint outputs and both only read Z
            main = torch.cuda.current_stream(self.device)
            self._fork.wait_stream(main)
            self._gram(Z, ks, n)
            with torch.cuda.stream(self._fork):
                self._loss(Z)
            main.wait_stream(self._fork)
            return
        self._loss(Z)
        self._gram(Z, ks, n)

    def supports_f16(self) -> bool:
        return True

    def host_tick(self):
        self._since += 1
        if self._pending is not None and self._since >= self._install_at:
            self._install()
        if self._since >= self.solve_every:
            if self.device.type == "cuda":
                self.solve_async()
            else:
                self.solve()

    def after_sync(self):
        pass   # the solve cadence is driven by `solveEvery`, and always runs before a query

    # ---- pipelined synchronisation (parallel/fused.py::PipelinedSynchronousDriver) -----------------
    def use_delta_buffers(self, bufs, drain) -> None:
        """Two symmetric delta buffers: updates go to the current one while the other is all-reduced."""
        ld = self.ldg
        self._deltas = [b.view(ld, ld) for b in bufs]
        self._drain = drain
        self.dG = self._deltas[0]

    def switch_delta(self, i: int) -> None:
        self.dG = self._deltas[i]

    def _snapshot(self) -> None:
        """_work <- global statistics as this rank knows them; must run on the stream that orders the Gram / sync
        kernels.  With the sharded hub this is the collective pull of the other owners' rows."""
        if self._hub is not None:
            self._hub.gather()
            self._work.copy_(self.G)
            return
        if self._deltas is not None:
            self._drain()                       # in-flight all-reduces update G and clear their delta
            torch.add(self.G, self._deltas[0], out=self._work)
            self._work.add_(self._deltas[1])
            return
        torch.add(self.G, self.dG, out=self._work)

    def _complete(self) -> None:
        """Upper blocks -> full symmetric matrix, then y_lo folded into y (row / column d + 2 into d + 1)."""
        d = self.d
        F.mirror_upper(self._work, d + 3)
        W = self._work
        W[: d + 3, d + 1] += W[: d + 3, d + 2]
        W[d + 1, : d + 3] += W[d + 2, : d + 3]
        W[: d + 3, d + 2] = 0
        W[d + 2, : d + 3] = 0

    def _solve_into(self, out: torch.Tensor, use_cg: bool = False) -> None:
        """out <- (G_A + lambda I)^-1 G_b from the completed _work: CG started from the current weights, or the
        blocked Cholesky kernel (which destroys _work)."""
        d1 = self.d + 1
        self._complete()
        if use_cg:
            if self._cg is None:
                self._cg = F.ridge_cg_workspace(d1, self.device, self.cg_iters)
            if out.data_ptr() != self.wt.data_ptr():
                out.copy_(self.wt[:d1])
            F.ridge_cg(self._work, d1, self.lam, out, *self._cg, max_iter=self.cg_iters, tol=self.cg_tol)
            return
        F.ridge_chol(self._work, d1, self.lam, out, self._chol_info)

    def solve_async(self):
        """Refresh w on a side stream so the refresh overlaps the next mini-batches; the result is installed a few
        steps later (`_install_at`)."""
        if self._pending is not None:
            self._install()
        use_cg = self.solver == "cg"
        if use_cg and self._cg_evt is not None and self._cg_evt.query():
            # verdict of the previous CG refresh, read without waiting: a refresh that did not reach the
            # tolerance is repeated by Cholesky now, three of them switch the solver for good
            self._cg_evt = None
            self.cg_last = (int(self._cg_host[0]), float(self._cg_host[1]))
            if not self.cg_last[1] <= 100.0 * self.cg_tol:
                self.cg_failures += 1
                use_cg = False
                if self.cg_failures >= 3:
                    self.solver = "cholesky"
        main = torch.cuda.current_stream(self.device)
        self._snapshot()                      # consistent copy, ordered with the Gram / sync kernels
        self._side.wait_stream(main)
        with torch.cuda.stream(self._side):
            self._solve_into(self._w_next, use_cg)
            if use_cg:
                if self._cg_host is None:
                    self._cg_host = torch.zeros(2, dtype=torch.float32).pin_memory()
                self._cg_host.copy_(self._cg[2], non_blocking=True)
                self._cg_evt = torch.cuda.Event()
                self._cg_evt.record(self._side)
            ev = torch.cuda.Event()
            ev.record(self._side)
        self._pending = ev
        self._since = 0
        self._install_at = min(8, max(1, self.solve_every - 1))
        self.solves += 1

    def _install(self):
        torch.cuda.current_stream(self.device).wait_event(self._pending)
        self.wt[: self.d + 1].copy_(self._w_next)
        self._pending = None

    def solve(self):
        """Synchronous refresh (queries, checkpoints, CPU).  Collective with the sharded hub."""
        if self._pending is not None:
            self._install()
        d1 = self.d + 1
        self._snapshot()
        ok = False
        if self.solver == "cg":
            self._solve_into(self.wt[:d1], True)
            it, rel = self._cg[2].tolist()
            self.cg_last = (int(it), rel)
            ok = rel <= 100.0 * self.cg_tol and bool(torch.isfinite(self.wt).all())
            if not ok:                                    # did not converge (or NaN): direct solve instead
                self.cg_failures += 1
        if not ok:
            w = torch.empty(d1, device=self.device)
            self._solve_into(w, False)                    # _work was only read by CG: complete() is idempotent
            self.wt[:d1].copy_(w)
            ok = float(self._chol_info[0]) == 0.0 and bool(torch.isfinite(self.wt).all())
        if not ok:                                        # not positive definite in fp32: redo in fp64
            self._snapshot_local_again()
            A = self._work[:d1, :d1].double()
            A.diagonal().add_(self.lam)
            try:
                self.wt[:d1] = torch.linalg.solve(A, self._work[:d1, d1].double()).float()
            except Exception:                             # singular (e.g. lambda = 0 before any data)
                self.wt[:d1] = torch.linalg.lstsq(A, self._work[:d1, d1].double()[:, None]).solution[:, 0].float()
        self._since = 0
        self.solves += 1

    def _snapshot_local_again(self) -> None:
        """Rebuild the completed _work after the Cholesky kernel destroyed it, WITHOUT another collective."""
        if self._hub is not None:
            self._work.copy_(self.G)
        elif self._deltas is not None:
            torch.add(self.G, self._deltas[0], out=self._work)
            self._work.add_(self._deltas[1])
        else:
            torch.add(self.G, self.dG, out=self._work)
        self._complete()

    def predict(self, Z):
        return Z[:, : self.d + 1] @ self.wt[: self.d + 1]

    def to_oracle(self):
        self.solve()
        self._snapshot_local_again()
        o = O.ORR(self.hyper)
        o.ensure_init(self.d)
        n = self.d + 2
        o.G = self._work[:n, :n].detach().double().cpu().numpy().copy()
        o.w = self.wt[: self.d + 1].detach().double().cpu().numpy().copy()
        c = self.counters()
        o.fitted, o.cum_loss = c["fitted"], c["loss_sum"]
        return o

    def from_oracle(self, o):
        n = self.d + 2
        self.G.zero_()
        self.dG.zero_()
        Gt = torch.from_numpy(o.G.astype(np.float32)).to(self.device)
        if self._hub is not None:               # running sums add up to the global matrix: rank 0 carries the warm start
            if self._hub.comm.rank == 0:
                self.dG[:n, :n] = Gt
            self._hub.push()
        else:
            self.G[:n, :n] = Gt
        self.solve()

    def state_dict(self):
        if self._deltas is not None:            # fold both local deltas into the saved one; G stays global
            self._drain()
            if self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
            sd = super().state_dict()
            sd["buf"] = (self._deltas[0] + self._deltas[1]).reshape(-1).detach().cpu().clone()
            return sd
        if self._hub is not None:               # buf = this worker's running sum, e = the gathered global statistics
            self._hub.gather()
            if self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
        return super().state_dict()

    def load_state_dict(self, sd):
        super().load_state_dict(sd)
        if self._deltas is not None:
            self._deltas[1].zero_()
        self.solve()
