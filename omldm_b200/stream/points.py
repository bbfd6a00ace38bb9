"""Data points flowing through the runtime (the reference's out-of-tree `mlAPI.math` types).

Call sites: utils/parsers/dataStream/DataPointParser.scala:20-46 -
`LabeledPoint(target, numerical, discrete, categorical, rawString)`,
`UnlabeledPoint(numerical, discrete, categorical, rawString)`, `TrainingPoint(point)`,
`ForecastingPoint(unlabeled)`; the raw JSON is kept for forecasts and nulled for training.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from ..api.pojos import DataInstance


@dataclass
class UsablePoint:
    numerical: np.ndarray
    discrete: np.ndarray = field(default_factory=lambda: np.zeros(0))
    categorical: List[str] = field(default_factory=list)
    data_instance: Optional[DataInstance] = None

    def features(self) -> np.ndarray:
        """numerical ++ discrete (ints widened to double, DataPointParser.scala:27-30)."""
        if self.discrete is None or len(self.discrete) == 0:
            return self.numerical
        return np.concatenate([self.numerical, self.discrete.astype(np.float64)])


@dataclass
class TrainingPoint(UsablePoint):
    target: float = 0.0


@dataclass
class ForecastingPoint(UsablePoint):
    pass


def data_point_from_instance(di: DataInstance) -> Optional[UsablePoint]:
    """DataPointParser.flatMap (DataPointParser.scala:16-55)."""
    if not di.is_valid():
        return None
    num = np.asarray(di.numericalFeatures or [], dtype=np.float64)
    disc = np.asarray(di.discreteFeatures or [], dtype=np.float64)
    cat = list(di.categoricalFeatures or [])
    feats = np.concatenate([num, disc]) if len(disc) else num
    if di.operation == "training":
        # unlabeled training points are legal (clustering); DataPointParser.scala:20-36
        return TrainingPoint(feats, disc, cat, None, float(di.target) if di.target is not None else 0.0)
    if di.operation == "forecasting":
        return ForecastingPoint(feats, disc, cat, di)
    return None


# ---- vector types of `mlAPI.math` that the reference registers for serialisation (CommonUtils.scala:17-23) -------------
class DenseVector:
    """Dense fp64 vector (`mlAPI.math.DenseVector`)."""

    def __init__(self, values):
        self.values = np.asarray(values, dtype=np.float64).reshape(-1)

    @property
    def size(self) -> int:
        return int(self.values.size)

    def to_dense(self) -> "DenseVector":
        return self

    def to_array(self) -> np.ndarray:
        return self.values

    def dot(self, other) -> float:
        return float(other.dot(self)) if isinstance(other, SparseVector) else float(self.values @ _arr(other))

    def __len__(self) -> int:
        return self.size

    def __eq__(self, o) -> bool:
        return isinstance(o, (DenseVector, SparseVector)) and np.array_equal(self.values, o.to_array())


class SparseVector:
    """Sparse fp64 vector (`mlAPI.math.SparseVector`): `size`, strictly increasing `indices`, `values`.  Duplicate
    indices given to the constructor are summed (the same rule the GPU densify / sparse kernels use)."""

    def __init__(self, size: int, indices, values):
        idx = np.asarray(indices, dtype=np.int64).reshape(-1)
        val = np.asarray(values, dtype=np.float64).reshape(-1)
        if idx.size != val.size:
            raise ValueError("indices and values differ in length")
        if idx.size and (idx.min() < 0 or idx.max() >= size):
            raise ValueError("index out of range")
        uniq, inv = np.unique(idx, return_inverse=True)
        self.size = int(size)
        self.indices = uniq
        self.values = np.bincount(inv, weights=val, minlength=uniq.size) if idx.size else val

    def to_array(self) -> np.ndarray:
        out = np.zeros(self.size)
        out[self.indices] = self.values
        return out

    def to_dense(self) -> DenseVector:
        return DenseVector(self.to_array())

    def dot(self, other) -> float:
        if isinstance(other, SparseVector):
            common, ia, ib = np.intersect1d(self.indices, other.indices, return_indices=True)
            return float(self.values[ia] @ other.values[ib])
        return float(self.values @ _arr(other)[self.indices])

    def to_ell(self, nnz: int):
        """(int32 indices[nnz], fp32 values[nnz]) padded with (0, 0.0): one record of a `SparseTopic` slot."""
        if self.indices.size > nnz:
            raise ValueError(f"{self.indices.size} non-zeros do not fit {nnz} slots")
        idx = np.zeros(nnz, dtype=np.int32)
        val = np.zeros(nnz, dtype=np.float32)
        idx[: self.indices.size] = self.indices
        val[: self.indices.size] = self.values
        return idx, val

    def __len__(self) -> int:
        return self.size

    def __eq__(self, o) -> bool:
        return isinstance(o, (DenseVector, SparseVector)) and np.array_equal(self.to_array(), o.to_array())


def _arr(v) -> np.ndarray:
    return v.to_array() if isinstance(v, (DenseVector, SparseVector)) else np.asarray(v, dtype=np.float64).reshape(-1)


class LabeledPoint(TrainingPoint):
    """`LabeledPoint(target, numerical, discrete, categorical, rawString)` (DataPointParser.scala:20-27): argument
    order of the reference, vectors accepted as arrays, DenseVector or SparseVector."""

    def __init__(self, target, numerical, discrete=None, categorical=None, raw: Optional[str] = None):
        super().__init__(_arr(numerical), _arr(discrete if discrete is not None else []), list(categorical or []),
                         None, float(target))
        self.raw = raw

    def as_unlabeled_point(self) -> "UnlabeledPoint":
        return UnlabeledPoint(self.numerical, self.discrete, self.categorical, self.raw)


class UnlabeledPoint(ForecastingPoint):
    """`UnlabeledPoint(numerical, discrete, categorical, rawString)` (DataPointParser.scala:28-36)."""

    def __init__(self, numerical, discrete=None, categorical=None, raw: Optional[str] = None):
        super().__init__(_arr(numerical), _arr(discrete if discrete is not None else []), list(categorical or []), None)
        self.raw = raw


# ---- remaining names of `mlAPI.math`: the reference's hierarchy is Point > {LearningPoint > {LabeledPoint,
# UnlabeledPoint}, UsablePoint > {TrainingPoint, ForecastingPoint}} with a UsablePoint WRAPPING a LearningPoint; here the
# wrapper and the wrapped point are one object, so the trait names are aliases of the common base.
Point = UsablePoint
LearningPoint = UsablePoint
Vector = (DenseVector, SparseVector)        # for isinstance checks: `isinstance(v, Vector)`
