"""fp64 oracle preprocessors: PolynomialFeatures, StandardScaler, MinMaxScaler
(whitelist utils/parsers/requestStream/PipelineMap.scala:67; maths out of tree).
"""
from __future__ import annotations

import itertools
from typing import Any, Dict, Optional

import numpy as np

from .learners import _h


def poly_factor_table(d: int, degree: int) -> np.ndarray:
    """[d_out, degree] int16 input indices per output monomial, -1 = unused.

    Highest degree first, no constant term, monomials of one degree in lexicographic order of
    their index multisets ((x, y), degree 2 -> x^2, xy, y^2, x, y): the FlinkML ordering the
    reference's library descends from (`registerFlinkMLTypes`, utils/CommonUtils.scala:14).
    """
    rows = []
    for deg in range(degree, 0, -1):
        for comb in itertools.combinations_with_replacement(range(d), deg):
            rows.append(list(comb) + [-1] * (degree - deg))
    return np.asarray(rows, dtype=np.int16).reshape(-1, degree)


class OraclePreprocessor:
    name = ""

    def __init__(self, hyper: Optional[dict] = None):
        self.hyper = dict(hyper or {})

    def out_dim(self, d: int) -> int:
        return d

    def fit_transform(self, x: np.ndarray) -> np.ndarray: ...
    def transform(self, x: np.ndarray) -> np.ndarray: ...

    def fit_transform_batch(self, X: np.ndarray) -> np.ndarray:
        return np.stack([self.fit_transform(x) for x in X]) if len(X) else X

    def transform_batch(self, X: np.ndarray) -> np.ndarray:
        return np.stack([self.transform(x) for x in X]) if len(X) else X

    def to_parameters(self) -> Dict[str, Any]:
        return {}


class PolynomialFeatures(OraclePreprocessor):
    name = "PolynomialFeatures"

    def __init__(self, hyper=None):
        super().__init__(hyper)
        self.degree = int(_h(hyper, ["degree"], 2))
        self._table = None

    def out_dim(self, d):
        return len(poly_factor_table(d, self.degree))

    def transform(self, x):
        if self._table is None or self._d != len(x):
            self._d = len(x)
            self._table = poly_factor_table(len(x), self.degree)
        xe = np.concatenate([np.asarray(x, dtype=np.float64), [1.0]])  # index -1 -> 1.0
        return np.prod(xe[self._table], axis=1)

    fit_transform = transform

    def transform_batch(self, X):
        return np.stack([self.transform(x) for x in X]) if len(X) else np.zeros((0, 0))

    fit_transform_batch = transform_batch


class StandardScaler(OraclePreprocessor):
    """Welford running mean / variance (sample variance, n-1); x -> (x - mean) / std."""
    name = "StandardScaler"

    def __init__(self, hyper=None):
        super().__init__(hyper)
        self.n = 0.0
        self.mean = None
        self.m2 = None

    def _ensure(self, d):
        if self.mean is None:
            self.mean, self.m2 = np.zeros(d), np.zeros(d)

    def fit(self, x):
        self._ensure(len(x))
        self.n += 1.0
        dlt = x - self.mean
        self.mean += dlt / self.n
        self.m2 += dlt * (x - self.mean)

    def fit_batch(self, X):
        """Chan et al. parallel merge - what the GPU kernel does per mini-batch."""
        if len(X) == 0:
            return
        self._ensure(X.shape[1])
        nb = float(len(X))
        mb = X.mean(0)
        m2b = ((X - mb) ** 2).sum(0)
        tot = self.n + nb
        delta = mb - self.mean
        self.m2 = self.m2 + m2b + delta ** 2 * self.n * nb / tot
        self.mean = self.mean + delta * nb / tot
        self.n = tot

    def transform(self, x):
        self._ensure(len(x))
        var = self.m2 / (self.n - 1.0) if self.n > 1 else np.zeros_like(self.m2)
        inv = np.where(var > 1e-24, 1.0 / np.sqrt(np.maximum(var, 1e-300)), 1.0)
        return (np.asarray(x, dtype=np.float64) - self.mean) * inv

    def fit_transform(self, x):
        x = np.asarray(x, dtype=np.float64)
        self.fit(x)
        return self.transform(x)

    def fit_transform_batch(self, X):
        X = np.asarray(X, dtype=np.float64)
        self.fit_batch(X)
        return self.transform_batch(X)

    def transform_batch(self, X):
        if len(X) == 0:
            return X
        self._ensure(X.shape[1])
        var = self.m2 / (self.n - 1.0) if self.n > 1 else np.zeros_like(self.m2)
        inv = np.where(var > 1e-24, 1.0 / np.sqrt(np.maximum(var, 1e-300)), 1.0)
        return (np.asarray(X, dtype=np.float64) - self.mean) * inv

    def to_parameters(self):
        return {"count": self.n, "mean": None if self.mean is None else self.mean.tolist(),
                "m2": None if self.m2 is None else self.m2.tolist()}


class MinMaxScaler(OraclePreprocessor):
    name = "MinMaxScaler"

    def __init__(self, hyper=None):
        super().__init__(hyper)
        self.mn = None
        self.mx = None

    def fit(self, x):
        if self.mn is None:
            self.mn, self.mx = np.array(x, dtype=np.float64), np.array(x, dtype=np.float64)
        else:
            self.mn, self.mx = np.minimum(self.mn, x), np.maximum(self.mx, x)

    def transform(self, x):
        x = np.asarray(x, dtype=np.float64)
        if self.mn is None:
            return x
        rng = self.mx - self.mn
        return (x - self.mn) * np.where(rng > 1e-30, 1.0 / np.maximum(rng, 1e-300), 0.0)

    def fit_transform(self, x):
        x = np.asarray(x, dtype=np.float64)
        self.fit(x)
        return self.transform(x)

    def fit_transform_batch(self, X):
        X = np.asarray(X, dtype=np.float64)
        if len(X):
            self.fit(X.min(0)); self.fit(X.max(0))
        return self.transform_batch(X)

    def transform_batch(self, X):
        return np.stack([self.transform(x) for x in X]) if len(X) else X

    def to_parameters(self):
        return {"min": None if self.mn is None else self.mn.tolist(),
                "max": None if self.mx is None else self.mx.tolist()}


ORACLE_PREPROCESSORS = {c.name: c for c in (PolynomialFeatures, StandardScaler, MinMaxScaler)}


def make_oracle_preprocessor(name: str, hyper: Optional[dict] = None) -> OraclePreprocessor:
    return ORACLE_PREPROCESSORS[name](hyper)
