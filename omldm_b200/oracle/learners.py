"""fp64 NumPy restatement of every learner: the correctness oracle and the timed CPU baseline.

The reference tree only names the learners (utils/parsers/requestStream/PipelineMap.scala:68);
their maths lives in the out-of-tree `mlAPI` project.  The formulas here are the literature
ones listed in SURVEY.md §8.2 and are the single source of truth the CUDA kernels are tested
against.  Conventions shared with the kernels:

* the intercept is a constant-1 feature appended to x (weight vector has d+1 entries);
* `fit` is prequential: it returns the loss of the prediction made BEFORE the update;
* `flat` is the vector the synchronisation protocols move; `merge_mode` says whether replicas
  combine by averaging ("mean") or by adding their deltas ("sum", additive sufficient statistics).
"""
from __future__ import annotations

import math
from typing import Any, Dict, List, Optional

import numpy as np

from ..api.compat import VectorBias


def _h(hyper: Optional[dict], keys, default):
    """Hyper-parameter lookup accepting several spellings ('accept common spellings')."""
    if hyper:
        for k in keys if isinstance(keys, (list, tuple)) else [keys]:
            if k in hyper and hyper[k] is not None:
                return hyper[k]
    return default


def pa_tau(loss: float, nrm2: float, variant: int, C: float) -> float:
    if loss <= 0.0 or nrm2 <= 0.0:
        return 0.0
    if variant == 0:
        return loss / nrm2
    if variant == 1:
        return min(C, loss / nrm2)
    return loss / (nrm2 + 1.0 / (2.0 * C))


def parse_variant(hyper: Optional[dict]) -> int:
    v = _h(hyper, ["updateType", "variant", "type"], "PA-I")
    if isinstance(v, (int, float)):
        return int(v)
    v = str(v).upper().replace("_", "-")
    return {"PA": 0, "PA-I": 1, "PA-1": 1, "PAI": 1, "PA-II": 2, "PA-2": 2, "PAII": 2}.get(v, 1)


class OracleLearner:
    name = ""
    task = "regression"
    merge_mode = "mean"

    def __init__(self, hyper: Optional[dict] = None):
        self.hyper = dict(hyper or {})
        self.d: Optional[int] = None
        self.fitted = 0
        self.cum_loss = 0.0
        self.mini_batch_update = str(_h(hyper, ["miniBatchUpdate"], "mean"))

    # -- to be provided by subclasses
    def _init(self, d: int) -> None: ...
    def predict(self, x: np.ndarray) -> float: ...
    def _fit(self, x: np.ndarray, y: float) -> float: ...
    def get_flat(self) -> np.ndarray: ...
    def set_flat(self, v: np.ndarray) -> None: ...
    def to_parameters(self) -> Dict[str, Any]: ...
    def from_parameters(self, p: Dict[str, Any]) -> None: ...

    def ensure_init(self, d: int) -> None:
        if self.d is None:
            self.d = int(d)
            self._init(self.d)

    def fit(self, x: np.ndarray, y: float) -> float:
        self.ensure_init(len(x))
        loss = float(self._fit(np.asarray(x, dtype=np.float64), float(y)))
        self.fitted += 1
        self.cum_loss += loss
        return loss

    def fit_batch(self, X: np.ndarray, y: np.ndarray) -> float:
        """Mini-batch semantics; default = record by record (exact)."""
        return sum(self.fit(X[i], y[i]) for i in range(len(X)))

    def predict_batch(self, X: np.ndarray) -> np.ndarray:
        return np.array([self.predict(x) for x in X])

    def score(self, X: np.ndarray, y: np.ndarray) -> float:
        """accuracy (classification), RMSE (regression), mean inertia (clustering)."""
        if len(X) == 0:
            return 0.0
        self.ensure_init(X.shape[1])
        p = self.predict_batch(X)
        if self.task == "classification":
            return float(np.mean(self._labels(p) == self._labels_true(y)))
        return float(math.sqrt(np.mean((p - y) ** 2)))

    def _labels(self, p):
        return np.where(p > 0, 1.0, -1.0)

    def _labels_true(self, y):
        return np.where(np.asarray(y) > 0, 1.0, -1.0)

    def hyper_parameters(self) -> Dict[str, Any]:
        return dict(self.hyper)


def _aug(x: np.ndarray) -> np.ndarray:
    return np.concatenate([x, [1.0]])


# ------------------------------------------------------------------------------ PA family
class PAClassifier(OracleLearner):
    name, task, merge_mode = "PA", "classification", "mean"

    def __init__(self, hyper=None):
        super().__init__(hyper)
        self.C = float(_h(hyper, ["C", "c"], 0.01))
        self.variant = parse_variant(hyper)

    def _init(self, d):
        self.w = np.zeros(d + 1)

    def predict(self, x):
        return float(self.w @ _aug(np.asarray(x, dtype=np.float64)))

    def _coef(self, f, y, nrm2):
        yy = 1.0 if y > 0 else -1.0
        loss = max(0.0, 1.0 - yy * f)
        return yy * pa_tau(loss, nrm2, self.variant, self.C), loss

    def _fit(self, x, y):
        z = _aug(x)
        c, loss = self._coef(float(self.w @ z), y, float(z @ z))
        self.w += c * z
        return loss

    def fit_batch(self, X, y):
        self.ensure_init(X.shape[1])
        Z = np.hstack([X, np.ones((len(X), 1))])
        f = Z @ self.w
        n2 = np.einsum("ij,ij->i", Z, Z)
        tot, delta, act = 0.0, np.zeros_like(self.w), 0
        for i in range(len(X)):
            c, loss = self._coef(f[i], y[i], n2[i])
            tot += loss
            if c != 0.0:
                delta += c * Z[i]
                act += 1
        if act:
            self.w += delta / act if self.mini_batch_update == "mean" else delta
        self.fitted += len(X)
        self.cum_loss += tot
        return tot

    def get_flat(self):
        return self.w

    def set_flat(self, v):
        self.w = np.array(v, dtype=np.float64)

    def to_parameters(self):
        return VectorBias.from_flat(self.w).to_parameters()

    def from_parameters(self, p):
        vb = VectorBias.from_parameters(p)       # (weights, bias) <-> the flat [w | b] layout
        self.d = len(vb.weights)
        self.w = vb.flat()


class PARegressor(PAClassifier):
    name, task = "RegressorPA", "regression"

    def __init__(self, hyper=None):
        super().__init__(hyper)
        self.eps = float(_h(hyper, ["epsilon", "eps"], 0.1))

    def _coef(self, f, y, nrm2):
        err = f - y
        loss = max(0.0, abs(err) - self.eps)
        return (-1.0 if err > 0 else 1.0) * pa_tau(loss, nrm2, self.variant, self.C), loss


class SVM(PAClassifier):
    """Pegasos: eta_t = 1/(lambda t); w <- (1 - eta lambda) w + eta y x [y w.x < 1]."""
    name, task = "SVM", "classification"

    def __init__(self, hyper=None):
        super().__init__(hyper)
        self.lam = float(_h(hyper, ["lambda", "lam", "regularization"], 0.01))
        self.t = 0

    def _fit(self, x, y):
        z = _aug(x)
        yy = 1.0 if y > 0 else -1.0
        f = float(self.w @ z)
        loss = max(0.0, 1.0 - yy * f)
        self.t += 1
        eta = 1.0 / (self.lam * self.t)
        self.w *= (1.0 - eta * self.lam)
        if yy * f < 1.0:
            self.w += eta * yy * z
        return loss

    def fit_batch(self, X, y):
        self.ensure_init(X.shape[1])
        Z = np.hstack([X, np.ones((len(X), 1))])
        yy = np.where(y > 0, 1.0, -1.0)
        f = Z @ self.w
        loss = np.maximum(0.0, 1.0 - yy * f)
        self.t += 1
        eta = 1.0 / (self.lam * self.t)
        viol = (yy * f < 1.0)
        self.w = (1.0 - eta * self.lam) * self.w + (eta / max(1, len(X))) * (Z[viol].T @ yy[viol])
        self.fitted += len(X)
        self.cum_loss += float(loss.sum())
        return float(loss.sum())


# ------------------------------------------------------------------------------ multiclass PA
class MultiClassPA(OracleLearner):
    name, task, merge_mode = "MultiClassPA", "classification", "mean"

    def __init__(self, hyper=None):
        super().__init__(hyper)
        self.C = float(_h(hyper, ["C", "c"], 0.01))
        self.variant = parse_variant(hyper)
        self.n_classes = int(_h(hyper, ["nClasses", "classes", "numClasses"], 10))

    def _init(self, d):
        self.W = np.zeros((self.n_classes, d + 1))

    def scores(self, x):
        return self.W @ _aug(np.asarray(x, dtype=np.float64))

    def predict(self, x):
        return float(np.argmax(self.scores(x)))

    def _coef(self, s, y, nrm2):
        y = int(round(y))
        s2 = s.copy()
        s2[y] = -np.inf
        r = int(np.argmax(s2))
        loss = max(0.0, 1.0 - (s[y] - s[r]))
        return y, r, pa_tau(loss, 2.0 * nrm2, self.variant, self.C), loss

    def _fit(self, x, y):
        z = _aug(x)
        yi, r, tau, loss = self._coef(self.W @ z, y, float(z @ z))
        if tau > 0:
            self.W[yi] += tau * z
            self.W[r] -= tau * z
        return loss

    def fit_batch(self, X, y):
        self.ensure_init(X.shape[1])
        Z = np.hstack([X, np.ones((len(X), 1))])
        S = Z @ self.W.T
        n2 = np.einsum("ij,ij->i", Z, Z)
        delta, act, tot = np.zeros_like(self.W), 0, 0.0
        for i in range(len(X)):
            yi, r, tau, loss = self._coef(S[i], y[i], n2[i])
            tot += loss
            if tau > 0:
                delta[yi] += tau * Z[i]
                delta[r] -= tau * Z[i]
                act += 1
        if act:
            self.W += delta / act if self.mini_batch_update == "mean" else delta
        self.fitted += len(X)
        self.cum_loss += tot
        return tot

    def _labels(self, p):
        return np.asarray(p)

    def _labels_true(self, y):
        return np.round(np.asarray(y))

    def get_flat(self):
        return self.W.reshape(-1)

    def set_flat(self, v):
        self.W = np.array(v, dtype=np.float64).reshape(self.n_classes, -1)

    def to_parameters(self):
        return {"a": self.W[:, :-1].reshape(-1).tolist(), "b": self.W[:, -1].tolist(),
                "nClasses": self.n_classes}

    def from_parameters(self, p):
        C = int(p.get("nClasses", self.n_classes))
        a = np.asarray(p["a"], dtype=np.float64).reshape(C, -1)
        b = np.atleast_1d(np.asarray(p["b"], dtype=np.float64)).reshape(C, 1)
        self.n_classes, self.d = C, a.shape[1]
        self.W = np.hstack([a, b])


# ------------------------------------------------------------------------------ ridge regression
class ORR(OracleLearner):
    """A = lambda I + sum z z^T, b = sum y z, w = A^-1 b with z = [x, 1].

    State is the augmented Gram matrix G = sum [z, y]^T [z, y]; (G, hence A - lambda I and b)
    is additive across replicas, so merge_mode is "sum".
    """
    name, task, merge_mode = "ORR", "regression", "sum"

    def __init__(self, hyper=None):
        super().__init__(hyper)
        self.lam = float(_h(hyper, ["lambda", "lam", "regularization"], 1.0))
        self.solve_every = int(_h(hyper, ["solveEvery"], 1))
        self._since_solve = 0

    def _init(self, d):
        self.G = np.zeros((d + 2, d + 2))
        self.w = np.zeros(d + 1)

    def solve(self):
        d1 = self.d + 1
        A = self.G[:d1, :d1] + self.lam * np.eye(d1)
        self.w = np.linalg.solve(A, self.G[:d1, d1])
        self._since_solve = 0
        return self.w

    def predict(self, x):
        return float(self.w @ _aug(np.asarray(x, dtype=np.float64)))

    def _fit(self, x, y):
        zy = np.concatenate([x, [1.0, y]])
        loss = (self.predict(x) - y) ** 2
        self.G += np.outer(zy, zy)
        self._since_solve += 1
        if self._since_solve >= self.solve_every:
            self.solve()
        return loss

    def fit_batch(self, X, y):
        self.ensure_init(X.shape[1])
        Zy = np.hstack([X, np.ones((len(X), 1)), np.asarray(y, dtype=np.float64)[:, None]])
        wt = np.concatenate([self.w, [-1.0]])
        loss = float(np.sum((Zy @ wt) ** 2))
        self.G += Zy.T @ Zy
        self.solve()
        self.fitted += len(X)
        self.cum_loss += loss
        return loss

    def get_flat(self):
        return self.G.reshape(-1)

    def set_flat(self, v):
        self.G = np.array(v, dtype=np.float64).reshape(self.d + 2, self.d + 2)
        self.solve()

    def to_parameters(self):
        d1 = self.d + 1
        A = self.G[:d1, :d1] + self.lam * np.eye(d1)
        return {"a": A.reshape(-1).tolist(), "b": self.G[:d1, d1].tolist(), "weights": self.w.tolist()}

    def from_parameters(self, p):
        b = np.atleast_1d(np.asarray(p["b"], dtype=np.float64))
        d1 = len(b)
        self.d = d1 - 1
        self.G = np.zeros((d1 + 1, d1 + 1))
        self.G[:d1, :d1] = np.asarray(p["a"], dtype=np.float64).reshape(d1, d1) - self.lam * np.eye(d1)
        self.G[:d1, d1] = b
        self.G[d1, :d1] = b
        self.solve()


# ------------------------------------------------------------------------------ k-means
class KMeans(OracleLearner):
    """MacQueen online k-means on additive statistics (S_j = sum of members, n_j = count)."""
    name, task, merge_mode = "K-means", "clustering", "sum"

    def __init__(self, hyper=None):
        super().__init__(hyper)
        self.k = int(_h(hyper, ["k", "K", "nClusters", "clusters"], 8))

    def _init(self, d):
        self.S = np.zeros((self.k, d + 1))  # last column = count

    def centroids(self):
        n = self.S[:, -1:]
        return np.divide(self.S[:, :-1], n, out=np.zeros_like(self.S[:, :-1]), where=n > 0)

    def _assign(self, x):
        n = self.S[:, -1]
        empty = np.where(n <= 0)[0]
        if len(empty):
            return int(empty[0]), 0.0
        d2 = np.sum((self.centroids() - x) ** 2, axis=1)
        j = int(np.argmin(d2))
        return j, float(d2[j])

    def predict(self, x):
        n = self.S[:, -1]
        if not np.any(n > 0):
            return 0.0
        d2 = np.sum((self.centroids() - np.asarray(x, dtype=np.float64)) ** 2, axis=1)
        d2[n <= 0] = np.inf
        return float(np.argmin(d2))

    def _fit(self, x, y):
        j, loss = self._assign(x)
        self.S[j, :-1] += x
        self.S[j, -1] += 1.0
        return loss

    def fit_batch(self, X, y=None):
        """Mini-batch: assign all against frozen centroids, then add members to the statistics."""
        self.ensure_init(X.shape[1])
        X = np.asarray(X, dtype=np.float64)
        n = self.S[:, -1]
        if not np.all(n > 0):  # bootstrap empty centroids record by record
            return sum(self.fit(X[i], 0.0) for i in range(len(X)))
        C = self.centroids()
        d2 = (X ** 2).sum(1)[:, None] - 2 * X @ C.T + (C ** 2).sum(1)[None, :]
        j = np.argmin(d2, axis=1)
        loss = float(np.maximum(d2[np.arange(len(X)), j], 0).sum())
        np.add.at(self.S[:, :-1], j, X)
        np.add.at(self.S[:, -1], j, 1.0)
        self.fitted += len(X)
        self.cum_loss += loss
        return loss

    def score(self, X, y=None):
        if len(X) == 0 or self.d is None:
            return 0.0
        C = self.centroids()[self.S[:, -1] > 0]
        if len(C) == 0:
            return 0.0
        d2 = ((np.asarray(X)[:, None, :] - C[None]) ** 2).sum(-1)
        return float(d2.min(1).mean())

    def get_flat(self):
        return self.S.reshape(-1)

    def set_flat(self, v):
        self.S = np.array(v, dtype=np.float64).reshape(self.k, -1)

    def to_parameters(self):
        return {"centroids": self.centroids().reshape(-1).tolist(), "counts": self.S[:, -1].tolist(), "k": self.k}

    def from_parameters(self, p):
        k = int(p.get("k", self.k))
        c = np.asarray(p["centroids"], dtype=np.float64).reshape(k, -1)
        n = np.atleast_1d(np.asarray(p["counts"], dtype=np.float64))
        self.k, self.d = k, c.shape[1]
        self.S = np.hstack([c * n[:, None], n[:, None]])


# ------------------------------------------------------------------------------ neural network
class NN(OracleLearner):
    """Small MLP trained by SGD; regression (MSE) or classification (softmax CE) by `nClasses`."""
    name, merge_mode = "NN", "mean"

    def __init__(self, hyper=None):
        super().__init__(hyper)
        self.hidden = [int(h) for h in _h(hyper, ["hiddenLayers", "hidden", "layers"], [32])]
        self.lr = float(_h(hyper, ["learningRate", "lr"], 0.01))
        self.n_classes = int(_h(hyper, ["nClasses", "classes"], 0))
        self.task = "classification" if self.n_classes > 1 else "regression"
        self.seed = int(_h(hyper, ["seed"], 0))

    def _init(self, d):
        rng = np.random.default_rng(self.seed)
        sizes = [d] + self.hidden + [self.n_classes if self.n_classes > 1 else 1]
        self.shapes = [(sizes[i + 1], sizes[i] + 1) for i in range(len(sizes) - 1)]
        self.layers = [rng.standard_normal(s) * math.sqrt(2.0 / s[1]) for s in self.shapes]
        for L in self.layers:
            L[:, -1] = 0.0

    def _forward(self, X):
        acts = [X]
        h = X
        for i, L in enumerate(self.layers):
            h = np.hstack([h, np.ones((len(h), 1))]) @ L.T
            if i < len(self.layers) - 1:
                h = np.maximum(h, 0.0)
            acts.append(h)
        return acts

    def predict_batch(self, X):
        out = self._forward(np.asarray(X, dtype=np.float64))[-1]
        return np.argmax(out, 1).astype(np.float64) if self.n_classes > 1 else out[:, 0]

    def predict(self, x):
        return float(self.predict_batch(np.asarray(x)[None])[0])

    def _fit(self, x, y):
        return self._step(x[None], np.array([y]))

    def fit_batch(self, X, y):
        self.ensure_init(X.shape[1])
        loss = self._step(np.asarray(X, dtype=np.float64), np.asarray(y, dtype=np.float64))
        self.fitted += len(X)
        self.cum_loss += loss
        return loss

    def _step(self, X, y):
        acts = self._forward(X)
        out = acts[-1]
        B = len(X)
        if self.n_classes > 1:
            m = out - out.max(1, keepdims=True)
            p = np.exp(m)
            p /= p.sum(1, keepdims=True)
            yi = np.round(y).astype(int)
            loss = float(-np.log(np.maximum(p[np.arange(B), yi], 1e-300)).sum())
            g = p
            g[np.arange(B), yi] -= 1.0
        else:
            err = out[:, 0] - y
            loss = float((err ** 2).sum())
            g = 2.0 * err[:, None]
        g = g / B
        for i in range(len(self.layers) - 1, -1, -1):
            a = np.hstack([acts[i], np.ones((B, 1))])
            gW = g.T @ a
            if i > 0:
                g = (g @ self.layers[i][:, :-1]) * (acts[i] > 0)
            self.layers[i] -= self.lr * gW
        return loss

    def _labels(self, p):
        return np.asarray(p) if self.n_classes > 1 else super()._labels(p)

    def _labels_true(self, y):
        return np.round(np.asarray(y)) if self.n_classes > 1 else super()._labels_true(y)

    def get_flat(self):
        return np.concatenate([L.reshape(-1) for L in self.layers])

    def set_flat(self, v):
        v = np.asarray(v, dtype=np.float64)
        o = 0
        for i, s in enumerate(self.shapes):
            n = s[0] * s[1]
            self.layers[i] = v[o:o + n].reshape(s).copy()
            o += n

    def to_parameters(self):
        return {f"layer{i}": L.reshape(-1).tolist() for i, L in enumerate(self.layers)}

    def from_parameters(self, p):
        for i, s in enumerate(self.shapes):
            self.layers[i] = np.asarray(p[f"layer{i}"], dtype=np.float64).reshape(s)


# ------------------------------------------------------------------------------ Hoeffding tree
class _Leaf:
    __slots__ = ("counts", "n", "mean", "m2", "since")

    def __init__(self, C, d):
        self.counts = np.zeros(C)
        self.n = np.zeros((C, d))
        self.mean = np.zeros((C, d))
        self.m2 = np.zeros((C, d))
        self.since = 0


class HoeffdingTree(OracleLearner):
    """VFDT with Gaussian numeric-attribute observers (Domingos & Hulten 2000)."""
    name, task, merge_mode = "HT", "classification", "none"

    def __init__(self, hyper=None):
        super().__init__(hyper)
        self.n_classes = int(_h(hyper, ["nClasses", "classes"], 2))
        self.grace = int(_h(hyper, ["gracePeriod", "nMin"], 200))
        self.delta = float(_h(hyper, ["delta", "splitConfidence"], 1e-7))
        self.tie = float(_h(hyper, ["tieThreshold", "tau"], 0.05))
        self.n_split_points = int(_h(hyper, ["splitPoints"], 10))
        self.max_nodes = int(_h(hyper, ["maxNodes"], 4096))

    def _init(self, d):
        # node arrays: feature (-1 = leaf), threshold, left, right
        self.feature: List[int] = [-1]
        self.threshold: List[float] = [0.0]
        self.left: List[int] = [-1]
        self.right: List[int] = [-1]
        self.leaves: Dict[int, _Leaf] = {0: _Leaf(self.n_classes, d)}

    def _cls(self, y):
        c = int(round(y))
        if self.n_classes == 2 and c < 0:
            c = 0
        return min(max(c, 0), self.n_classes - 1)

    def _route(self, x):
        i = 0
        while self.feature[i] >= 0:
            i = self.left[i] if x[self.feature[i]] <= self.threshold[i] else self.right[i]
        return i

    def predict(self, x):
        leaf = self.leaves[self._route(np.asarray(x, dtype=np.float64))]
        return float(np.argmax(leaf.counts)) if leaf.counts.sum() > 0 else 0.0

    def _labels(self, p):
        return np.asarray(p)

    def _labels_true(self, y):
        return np.array([self._cls(v) for v in np.asarray(y)], dtype=np.float64)

    @staticmethod
    def _entropy(c):
        t = c.sum()
        if t <= 0:
            return 0.0
        p = c[c > 0] / t
        return float(-(p * np.log2(p)).sum())

    def _fit(self, x, y):
        c = self._cls(y)
        i = self._route(x)
        leaf = self.leaves[i]
        loss = 0.0 if (leaf.counts.sum() > 0 and int(np.argmax(leaf.counts)) == c) else 1.0
        leaf.counts[c] += 1
        leaf.n[c] += 1
        dlt = x - leaf.mean[c]
        leaf.mean[c] += dlt / leaf.n[c]
        leaf.m2[c] += dlt * (x - leaf.mean[c])
        leaf.since += 1
        if leaf.since >= self.grace and len(self.feature) + 2 <= self.max_nodes:
            leaf.since = 0
            self._try_split(i, leaf)
        return loss

    def _try_split(self, i, leaf):
        from math import erf, sqrt, log
        tot = leaf.counts.sum()
        if tot <= 0 or np.count_nonzero(leaf.counts) < 2:
            return
        h0 = self._entropy(leaf.counts)
        best = []
        for f in range(self.d):
            present = leaf.n[:, f] > 0
            if not present.any():
                continue
            mu, n = leaf.mean[:, f], leaf.n[:, f]
            sd = np.sqrt(np.divide(leaf.m2[:, f], np.maximum(n - 1, 1)))
            lo, hi = (mu - 2 * sd)[present].min(), (mu + 2 * sd)[present].max()
            if hi <= lo:
                continue
            gbest, tbest = -1.0, lo
            for t in np.linspace(lo, hi, self.n_split_points + 2)[1:-1]:
                cdf = np.array([0.5 * (1 + erf((t - mu[k]) / (sd[k] * sqrt(2)))) if sd[k] > 0 else float(mu[k] <= t)
                                for k in range(self.n_classes)])
                left = leaf.counts * cdf
                right = leaf.counts - left
                g = h0 - (left.sum() * self._entropy(left) + right.sum() * self._entropy(right)) / tot
                if g > gbest:
                    gbest, tbest = g, t
            best.append((gbest, f, tbest))
        if not best:
            return
        best.sort(reverse=True)
        g1 = best[0][0]
        g2 = best[1][0] if len(best) > 1 else 0.0
        R = math.log2(max(2, self.n_classes))
        eps = math.sqrt(R * R * log(1.0 / self.delta) / (2.0 * tot))
        if g1 > 0 and (g1 - g2 > eps or eps < self.tie):
            _, f, t = best[0]
            li, ri = len(self.feature), len(self.feature) + 1
            self.feature[i], self.threshold[i], self.left[i], self.right[i] = f, float(t), li, ri
            for _ in range(2):
                self.feature.append(-1); self.threshold.append(0.0); self.left.append(-1); self.right.append(-1)
            del self.leaves[i]
            for j in (li, ri):
                self.leaves[j] = _Leaf(self.n_classes, self.d)
            # children start with the parent's class prior split by the Gaussian estimate
            for k in range(self.n_classes):
                n, mu = leaf.n[k, f], leaf.mean[k, f]
                sd = math.sqrt(leaf.m2[k, f] / max(n - 1, 1)) if n > 1 else 0.0
                cdf = 0.5 * (1 + erf((t - mu) / (sd * sqrt(2)))) if sd > 0 else float(mu <= t)
                self.leaves[li].counts[k] = leaf.counts[k] * cdf
                self.leaves[ri].counts[k] = leaf.counts[k] * (1 - cdf)

    def get_flat(self):
        return np.zeros(0)

    def set_flat(self, v):
        pass

    def to_parameters(self):
        return {}

    def data_structure(self):
        leaves = {str(i): {"counts": l.counts.tolist()} for i, l in self.leaves.items()}
        return {"feature": list(self.feature), "threshold": list(self.threshold), "left": list(self.left),
                "right": list(self.right), "leaves": leaves}

    def from_parameters(self, p):
        pass


ORACLE_LEARNERS = {c.name: c for c in (PAClassifier, PARegressor, ORR, SVM, MultiClassPA, KMeans, NN, HoeffdingTree)}


def make_oracle_learner(name: str, hyper: Optional[dict] = None) -> OracleLearner:
    if name not in ORACLE_LEARNERS:
        raise KeyError(f"unknown learner {name!r}")
    return ORACLE_LEARNERS[name](hyper)
