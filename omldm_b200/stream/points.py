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
