"""Device learners, part 2: k-means, multiclass PA, neural network, Hoeffding tree (see learners.py)."""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from ..ops import functional as F
from ..oracle import learners as O
from ..oracle.learners import _h, parse_variant
from .learners import GpuLearner, GpuORR, GpuPA, GpuRegressorPA, GpuSVM, pad4


class GpuKMeans(GpuLearner):
    """Online k-means on additive statistics; distances X C^T and updates T^T Z on tcgen05.

    Distributed k-means (BASELINE.json config 4) is a capability the reference does not have: it
    forces K-means to a single central learner (FlinkSpoke.scala:207-209).
    """
    name, task, merge_mode = "K-means", "clustering", "sum"

    def __init__(self, hyper, device):
        super().__init__(hyper, device)
        self.k = int(_h(hyper, ["k", "K", "nClusters", "clusters"], 8))
        self._boot = False

    def n_flat(self, d):
        return self.k * F.pad32(d + 2)

    def _init(self):
        K, ld = self.k, self.ld
        self.dS = self.buf.view(K, ld)
        self.S = self.e.view(K, ld)
        self.centroids = torch.zeros(K, ld, device=self.device)
        self.cnorm = torch.full((K,), float("inf"), device=self.device)
        self._T = None
        self._Sc = None

    def _bufs(self, B):
        if self._T is None or self._T.shape[0] < B:
            self._T = torch.empty(B, F.pad32(self.k), device=self.device)
            self._Sc = torch.empty(B, pad4(self.k), device=self.device)
        return self._T[:B], self._Sc[:B]

    def refresh(self):
        F.kmeans_finalize(self.dS, self.S, self.k, self.d, self.centroids, self.cnorm)

    def graph_ready(self):
        return self._boot and self.mode != "exact"

    def after_sync(self):
        self.refresh()

    def fit_batch(self, Z):
        B = Z.shape[0]
        start = 0
        if not self._boot:
            # bootstrap: the first records claim the empty centroids in exact (MacQueen) order
            nb = min(B, max(2 * self.k, 64))
            F.kmeans_exact(Z[:nb], self.d, self.dS, self.k, self.stats)
            self.refresh()
            self._boot = True
            start = nb
            if start >= B:
                return
        Zb = Z[start:]
        if self.mode == "exact":
            F.kmeans_exact(Zb, self.d, self.dS, self.k, self.stats)   # against local statistics only
            self.refresh()
            return
        T, Sc = self._bufs(Zb.shape[0])
        # distances, argmin and the one-hot rows in one tcgen05 kernel (argmin read from TMEM); two kernels otherwise
        if not F.kmeans_assign_fused(Zb, self.d, self.centroids, self.cnorm, self.k, T, self.stats):
            F.gemm_tn(Zb[:, : self.d], self.centroids[:, : self.d], Sc[:, : self.k])
            F.kmeans_assign(Zb, self.d, Sc, self.cnorm, self.k, T, self.stats)
        F.gemm_outer(T, Zb, self.dS, accumulate=True, m=self.k, n=self.d + 1,
                     ksplit=max(1, min(32, Zb.shape[0] // 256)))
        self.refresh()

    def predict(self, Z):
        T, Sc = self._bufs(Z.shape[0])
        F.gemm_tn(Z[:, : self.d], self.centroids[:, : self.d], Sc[:, : self.k])
        return (self.cnorm[None, :] - 2 * Sc[:, : self.k]).argmin(1).float()

    def score(self, Z):
        if Z.shape[0] == 0 or not self._boot:
            return 0.0
        valid = Z[:, self.d] != 0
        if not bool(valid.any()):
            return 0.0
        T, Sc = self._bufs(Z.shape[0])
        F.gemm_tn(Z[:, : self.d], self.centroids[:, : self.d], Sc[:, : self.k])
        dist = (Z[:, : self.d] ** 2).sum(1) + (self.cnorm[None, :] - 2 * Sc[:, : self.k]).min(1).values
        return float(dist[valid].clamp_min(0).mean().item())

    def to_oracle(self):
        o = O.KMeans(self.hyper)
        o.ensure_init(self.d)
        tot = (self.S + self.dS)[:, : self.d + 1].detach().double().cpu().numpy()
        o.S = tot.copy()
        c = self.counters()
        o.fitted, o.cum_loss = c["fitted"], c["loss_sum"]
        return o

    def from_oracle(self, o):
        self.S.zero_(); self.dS.zero_()
        self.S[:, : self.d + 1] = torch.from_numpy(o.S.astype(np.float32)).to(self.device)
        self._boot = bool((o.S[:, -1] > 0).any())
        self.refresh()

    def load_state_dict(self, sd):
        super().load_state_dict(sd)
        self._boot = bool(((self.S + self.dS)[:, self.d] > 0).any().item())


class GpuMultiClassPA(GpuLearner):
    """Multiclass PA: scores X W^T and the update T^T X both on tcgen05 (csrc/umma_gemm.cu)."""
    name, task, merge_mode = "MultiClassPA", "classification", "mean"

    def __init__(self, hyper, device):
        super().__init__(hyper, device)
        self.C = float(_h(hyper, ["C", "c"], 0.01))
        self.variant = parse_variant(hyper)
        self.n_classes = int(_h(hyper, ["nClasses", "classes", "numClasses"], 10))
        self.mean_update = str(_h(hyper, ["miniBatchUpdate"], "mean")) == "mean"
        self.max_rows = int(_h(hyper, ["exactChunk"], 1))

    def n_flat(self, d):
        return self.n_classes * F.pad32(d + 2)

    def _init(self):
        self.W = self.buf.view(self.n_classes, self.ld)
        self._T = None
        self._coef = None
        self._wt = None
        self._act = torch.zeros(1, device=self.device)

    def _bufs(self, B):
        if self._T is None or self._T.shape[0] < B:
            self._T = torch.empty(B, F.pad32(self.n_classes), device=self.device)
            self._S = torch.empty(B, pad4(self.n_classes), device=self.device)
        return self._T[:B], self._S[:B]

    def _step(self, Z, sqnorm=None):
        C, d = self.n_classes, self.d
        T, S = self._bufs(Z.shape[0])
        ks = max(1, min(16, (d + 1) // 4096))
        F.gemm_tn(Z[:, : d + 1], self.W[:, : d + 1], S[:, :C], ksplit=ks)
        if sqnorm is None:
            sqnorm = F.row_sqnorm(Z, self.ld)   # pad columns are zero, target excluded below
            sqnorm = sqnorm - Z[:, d + 1] ** 2
        prev = self.stats[3].clone()
        F.mcpa_coef(Z, d, S, C, sqnorm, self.variant, self.C, T, self.stats)
        if self.mean_update:
            self._act.copy_((self.stats[3] - prev).float().reshape(1))
            F.scale_rows(T, C, self._act, reciprocal=True)
        F.gemm_outer(T, Z, self.W, accumulate=True, m=C, n=d + 1, ksplit=max(1, min(8, Z.shape[0] // 256)))

    def graph_ready(self):
        return self.mode != "exact"

    def supports_sparse(self) -> bool:
        return self.mode != "exact" and 2 <= self.n_classes <= 32

    def fit_sparse(self, rows, nnz):
        """Mini-batch step straight from sparse rows (csrc/sparse_mcpa.cu): gathers and scatters against the
        dense model instead of two GEMMs over a rebuilt dense batch."""
        if F.native(rows):
            B = rows.shape[0]
            if self._coef is None or self._coef.shape[0] < B:
                self._coef = torch.empty(B, 4, device=self.device)
            if self._wt is None and self.n_classes <= 16:
                self._wt = torch.empty((self.d + 1) * 16, device=self.device)
            prev = self.stats[3].clone()
            F.mcpa_sparse_step(rows, nnz, self.W, self.d, self.n_classes, self.variant, self.C, self._coef,
                               self.stats, prev, self.mean_update, self._wt)
            return
        Z = F.sparse_rows_to_dense(rows, nnz, self.d)
        col = rows[:, :nnz].contiguous().view(torch.int32)
        v = rows[:, nnz:2 * nnz]
        sq = torch.where((col >= 0) & (col < self.d), v * v, torch.zeros_like(v)).sum(1) + 1.0
        self._step(Z, sq)

    def fit_batch(self, Z, sqnorm=None):
        if self.mode == "exact":
            for r in range(Z.shape[0]):   # per-record order; the GEMMs degenerate to GEMVs
                self._step(Z[r:r + 1], None if sqnorm is None else sqnorm[r:r + 1])
        else:
            self._step(Z, sqnorm)

    def predict(self, Z):
        T, S = self._bufs(Z.shape[0])
        F.gemm_tn(Z[:, : self.d + 1], self.W[:, : self.d + 1], S[:, : self.n_classes])
        return S[:, : self.n_classes].argmax(1).float()

    def _labels(self, p):
        return p

    def _labels_true(self, y):
        return torch.round(y)

    def to_oracle(self):
        o = O.MultiClassPA(self.hyper)
        o.ensure_init(self.d)
        o.W = self.W[:, : self.d + 1].detach().double().cpu().numpy().copy()
        c = self.counters()
        o.fitted, o.cum_loss = c["fitted"], c["loss_sum"]
        return o

    def from_oracle(self, o):
        self.W.zero_()
        self.W[:, : self.d + 1] = torch.from_numpy(o.W.astype(np.float32)).to(self.device)


class GpuNN(GpuLearner):
    """Small MLP (ReLU) trained by mini-batch SGD; forward X W^T and weight gradients dH^T A run on
    the tcgen05 GEMMs, the element-wise steps between them (activation, output loss / gradient, relu mask, SGD
    step with the batch normalisation read from device memory) are fused kernels (csrc/nn.cu)."""
    name, merge_mode = "NN", "mean"

    def __init__(self, hyper, device):
        super().__init__(hyper, device)
        self.hidden = [int(h) for h in _h(hyper, ["hiddenLayers", "hidden", "layers"], [32])]
        self.lr = float(_h(hyper, ["learningRate", "lr"], 0.01))
        self.n_classes = int(_h(hyper, ["nClasses", "classes"], 0))
        self.task = "classification" if self.n_classes > 1 else "regression"
        self.seed = int(_h(hyper, ["seed"], 0))

    def _sizes(self, d):
        return [d] + self.hidden + [self.n_classes if self.n_classes > 1 else 1]

    def n_flat(self, d):
        s = self._sizes(d)
        return sum(s[i + 1] * F.pad32(s[i] + 1) for i in range(len(s) - 1))

    def _init(self):
        s = self._sizes(self.d)
        ref = O.NN(self.hyper)
        ref.ensure_init(self.d)       # same seed -> same initial weights as the oracle, on every rank
        self.layers = []
        o = 0
        for i in range(len(s) - 1):
            ldw = F.pad32(s[i] + 1)
            W = self.buf[o:o + s[i + 1] * ldw].view(s[i + 1], ldw)
            W.zero_()
            W[:, : s[i] + 1] = torch.from_numpy(ref.layers[i].astype(np.float32)).to(self.device)
            self.layers.append(W)
            o += s[i + 1] * ldw
        self.sizes = s
        self._bufB = {}
        self._gW = None
        if self.e is not None:
            self.e.copy_(self.buf)

    def _buffers(self, B):
        """Per-batch-size work buffers (activations, gradient rows, weight gradients); a few sizes are kept (training
        batch, holdout, forecast) so alternating between them does not re-allocate."""
        if not isinstance(self._bufB, dict):
            self._bufB = {}
        b = self._bufB.get(B)
        if b is None:
            if len(self._bufB) >= 4:
                self._bufB.pop(next(iter(self._bufB)))
            dev = self.device
            b = ([torch.zeros(B, F.pad32(self.sizes[i + 1] + 1), device=dev) for i in range(len(self.layers))],
                 [torch.zeros(B, F.pad32(self.sizes[i + 1]), device=dev) for i in range(len(self.layers))])
            self._bufB[B] = b
        if self._gW is None:
            self._gW = [torch.zeros_like(W) for W in self.layers]
            self._nv = torch.zeros(1, device=self.device)
        return b[0], b[1], self._gW

    def _forward(self, Z):
        """Returns activations with a trailing ones column (pitch pad32); GEMMs on tcgen05, activation fused
        (csrc/nn.cu::nn_hidden_act_kernel)."""
        hs, _, _ = self._buffers(Z.shape[0])
        acts = [Z]
        a, fan_in = Z, self.d
        for i, W in enumerate(self.layers):
            out_dim = self.sizes[i + 1]
            H = hs[i]
            F.gemm_tn(a[:, : fan_in + 1], W[:, : fan_in + 1], H[:, :out_dim])
            if i < len(self.layers) - 1:
                F.nn_hidden_act(H, out_dim)
            acts.append(H)
            a, fan_in = H, out_dim
        return acts

    def fit_batch(self, Z):
        d = self.d
        acts = self._forward(Z)
        _, Gs, gWs = self._buffers(Z.shape[0])
        last = len(self.layers) - 1
        self._nv.zero_()
        # loss statistics + dL/d(out) for the valid rows; the 1 / n_valid of the mean goes into the SGD step
        F.nn_out_grad(acts[-1], Z, d, self.sizes[-1], self.n_classes, Gs[last], self.stats, self._nv)
        for i in range(last, -1, -1):
            fan_in, od = self.sizes[i], self.sizes[i + 1]
            W, G = self.layers[i], Gs[i]
            if i > 0:
                # gradient w.r.t. the previous activations, through the relu: G (K-major) times W (rows = K, MN-major)
                F.gemm_nn(G[:, :od], W, Gs[i - 1][:, :fan_in])
                F.nn_relu_mask(Gs[i - 1], acts[i], fan_in)
            gW = gWs[i]
            gW.zero_()
            F.gemm_outer(G, acts[i], gW, accumulate=True, m=od, n=fan_in + 1)
            F.nn_sgd_step(W, gW, self.lr, self._nv)

    def predict(self, Z):
        out = self._forward(Z)[-1][:, : self.sizes[-1]]
        return out.argmax(1).float() if self.n_classes > 1 else out[:, 0]

    def _labels(self, p):
        return p if self.n_classes > 1 else super()._labels(p)

    def _labels_true(self, y):
        return torch.round(y) if self.n_classes > 1 else super()._labels_true(y)

    def to_oracle(self):
        o = O.NN(self.hyper)
        o.ensure_init(self.d)
        for i, W in enumerate(self.layers):
            o.layers[i] = W[:, : self.sizes[i] + 1].detach().double().cpu().numpy().copy()
        c = self.counters()
        o.fitted, o.cum_loss = c["fitted"], c["loss_sum"]
        return o

    def from_oracle(self, o):
        for i, W in enumerate(self.layers):
            W.zero_()
            W[:, : self.sizes[i] + 1] = torch.from_numpy(o.layers[i].astype(np.float32)).to(self.device)


class GpuHT(GpuLearner):
    """Hoeffding tree.  The reference never distributes it (forced SingleLearner,
    FlinkSpoke.scala:204-206).  Routing and the per-leaf sufficient statistics of the Gaussian
    attribute observers run on the GPU (csrc/tree.cu); the small, irregular tree and the Hoeffding-bound
    split test stay on the host and are evaluated per mini-batch for the leaves that passed their grace
    period.  `trainingMode: "exact"` runs the fp64 per-record oracle on the host instead."""
    name, task, merge_mode = "HT", "classification", "none"

    def __init__(self, hyper, device):
        super().__init__(hyper, device)
        self.tree = O.HoeffdingTree(hyper)          # host mirror: structure + split logic

    def n_flat(self, d):
        return 4

    def needs_e(self):
        return False

    def graph_ready(self):
        return False                                 # split decisions need a host round trip per batch

    def _init(self):
        t = self.tree
        t.ensure_init(self.d)
        C, d = t.n_classes, self.d
        cap = max(3, min(t.max_nodes, int(256e6 // (8 * C * d))))   # statistics stay below ~256 MB
        cap -= (cap + 1) % 2                                         # odd: root + pairs of children
        self.cap = t.max_nodes = cap
        dev = self.device
        self.feature = torch.full((cap,), -1, dtype=torch.int32, device=dev)
        self.threshold = torch.zeros(cap, device=dev)
        self.left = torch.full((cap,), -1, dtype=torch.int32, device=dev)
        self.right = torch.full((cap,), -1, dtype=torch.int32, device=dev)
        self.slot = torch.arange(cap, dtype=torch.int32, device=dev)  # statistics slot = node id
        self.cnt = torch.zeros(cap, C, device=dev)
        self.sum = torch.zeros(cap, C, d, device=dev)
        self.sq = torch.zeros(cap, C, d, device=dev)
        self.checked = {0: 0.0}
        # on-device split test (csrc/tree.cu::omldm_ht_split): the tree then grows without a host round trip and the
        # host mirror is refreshed lazily (`_pull_tree`)
        self.split_on_device = (str(_h(self.hyper, ["splitOnDevice"], "true")).lower() in ("1", "true", "yes")
                                and 2 <= C <= 16)
        self.n_nodes_dev = torch.ones(1, dtype=torch.int32, device=dev)
        self.checked_dev = torch.zeros(cap, device=dev)
        self._decision = torch.zeros(cap, 4, dtype=torch.int32, device=dev)
        self._n_splits = torch.zeros(1, dtype=torch.int32, device=dev)
        self._mirror_stale = False

    # ---- training ---------------------------------------------------------------------------------------
    def fit_batch(self, Z):
        if self.mode == "exact":
            h = Z[:, : self.d + 2].detach().cpu().double().numpy()
            for r in h:
                if r[self.d] != 0:
                    self.tree.fit(r[: self.d], r[self.d + 1])
            self.stats[0] = self.tree.cum_loss
            self.stats[2] = self.tree.fitted
            return
        t = self.tree
        F.ht_accumulate(Z, self.d, self.feature, self.threshold, self.left, self.right, self.slot, t.n_classes,
                        self.cnt, self.sum, self.sq, self.stats)
        self._maybe_split()

    def _pull_tree(self) -> None:
        """Refresh the host mirror (structure + leaf class counts) from the device arrays."""
        if not self._mirror_stale:
            return
        t = self.tree
        n = int(self.n_nodes_dev.item())
        t.feature = self.feature[:n].tolist()
        t.threshold = [float(v) for v in self.threshold[:n].tolist()]
        t.left, t.right = self.left[:n].tolist(), self.right[:n].tolist()
        cnt_h = self.cnt[:n].detach().cpu().double().numpy()
        chk = self.checked_dev[:n].tolist()
        t.leaves = {}
        self.checked = {}
        for i in range(n):
            if t.feature[i] < 0:
                lf = O._Leaf(t.n_classes, self.d)
                lf.counts = cnt_h[i].copy()
                t.leaves[i] = lf
                self.checked[i] = float(chk[i])
        self._mirror_stale = False

    def _maybe_split(self):
        t = self.tree
        if self.split_on_device and F.native(self.cnt):
            F.ht_split(self.feature, self.threshold, self.left, self.right, self.n_nodes_dev, t.n_classes, self.d,
                       self.cnt, self.sum, self.sq, self.checked_dev, float(t.grace), float(t.delta), float(t.tie),
                       int(t.n_split_points), self._decision, self._n_splits)
            self._mirror_stale = True
            return
        n_nodes = len(t.feature)
        cnt_h = self.cnt[:n_nodes].detach().cpu().double().numpy()
        for leaf in [i for i in range(n_nodes) if t.feature[i] < 0]:
            tot = float(cnt_h[leaf].sum())
            if tot - self.checked.get(leaf, 0.0) < t.grace or len(t.feature) + 2 > self.cap:
                continue
            self.checked[leaf] = tot
            lf = O._Leaf(t.n_classes, self.d)
            lf.counts = cnt_h[leaf].copy()
            s = self.sum[leaf].detach().cpu().double().numpy()
            q = self.sq[leaf].detach().cpu().double().numpy()
            n = np.maximum(lf.counts[:, None], 0.0) * np.ones((1, self.d))
            lf.n = n
            lf.mean = np.divide(s, n, out=np.zeros_like(s), where=n > 0)
            lf.m2 = np.maximum(q - n * lf.mean ** 2, 0.0)
            t.leaves[leaf] = lf
            before = len(t.feature)
            t._try_split(leaf, lf)
            if len(t.feature) > before:                # split happened: publish it to the device arrays
                li, ri = t.left[leaf], t.right[leaf]
                self.feature[leaf] = t.feature[leaf]
                self.threshold[leaf] = float(t.threshold[leaf])
                self.left[leaf], self.right[leaf] = li, ri
                for ch in (li, ri):                    # children start from the parent's class prior
                    self.cnt[ch] = torch.from_numpy(t.leaves[ch].counts.astype(np.float32)).to(self.device)
                    self.checked[ch] = float(t.leaves[ch].counts.sum())
        t.fitted = int(self.stats[2].item())
        t.cum_loss = float(self.stats[0].item())

    # ---- inference -------------------------------------------------------------------------------------------
    def predict(self, Z):
        if self.mode == "exact":
            h = Z[:, : self.d].detach().cpu().double().numpy()
            return torch.tensor([self.tree.predict(r) for r in h], device=self.device, dtype=torch.float32)
        leaf = F.ht_route(Z, self.feature, self.threshold, self.left, self.right)
        return self.cnt[leaf].argmax(1).float()

    def _labels(self, p):
        return p

    def _labels_true(self, y):
        return F.ht_class(y, self.tree.n_classes).float()

    def to_oracle(self):
        t = self.tree
        if self.mode != "exact":
            self._pull_tree()
            t.fitted = int(self.stats[2].item())
            t.cum_loss = float(self.stats[0].item())
            cnt_h = self.cnt[: len(t.feature)].detach().cpu().double().numpy()
            for i in range(len(t.feature)):
                if t.feature[i] < 0:
                    lf = t.leaves.get(i) or O._Leaf(t.n_classes, self.d)
                    lf.counts = cnt_h[i].copy()
                    t.leaves[i] = lf
        return t

    def from_oracle(self, o):
        self.tree = o

    def data_structure(self):
        return self.to_oracle().data_structure()

    def state_dict(self):
        sd = super().state_dict()
        sd["tree"] = self.to_oracle()
        if self.mode != "exact":
            sd["ht"] = {k: getattr(self, k).detach().cpu().clone() for k in
                        ("feature", "threshold", "left", "right", "cnt", "sum", "sq", "n_nodes_dev", "checked_dev")}
            sd["checked"] = dict(self.checked)
        return sd

    def load_state_dict(self, sd):
        self.stats.copy_(sd["stats"].to(self.device))
        self.tree = sd["tree"]
        if "ht" in sd:
            for k, v in sd["ht"].items():
                getattr(self, k).copy_(v.to(self.device))
            self.checked = dict(sd["checked"])
            if "n_nodes_dev" not in sd["ht"]:          # checkpoint written before the on-device split test
                self.n_nodes_dev.fill_(len(self.tree.feature))
                for i, v in self.checked.items():
                    self.checked_dev[int(i)] = float(v)


GPU_LEARNERS = {c.name: c for c in (GpuPA, GpuRegressorPA, GpuSVM, GpuORR, GpuKMeans, GpuMultiClassPA, GpuNN, GpuHT)}


def make_gpu_learner(name: str, hyper: Optional[dict], device) -> GpuLearner:
    if name not in GPU_LEARNERS:
        raise KeyError(f"unknown learner {name!r}")
    return GPU_LEARNERS[name](hyper, device)
