"""Small pieces of the out-of-tree API surface that the reference job touches by name, re-supplied so
that code written against those names has something to hold on to.

* `CountableSerial`          - anything with `get_size()` (FlinkMessage.scala:8,16-23; Job.scala:32-33): POJOs, messages.
* `Mergeable`                - anything with `merge(others)` (SpokeLogic.scala:41-47): DataSet, wrappers, learner states.
* `LearningParameters`, `VectorBias(weights, bias)` - linear-model parameters are a dense fp64 vector plus a scalar
  intercept (StateAccumulators.scala:8,25-27); here a thin view over the flat `[w | b]` vector the learners keep.
* `integer_parsing` / `double_parsing` - typed lookups in a configuration map with a default
  (`utils.Parsing.IntegerParsing`, FlinkSpoke.scala:184, FlinkHub.scala:177).
"""
from __future__ import annotations

from typing import Any, Dict, Iterable, Optional, Protocol, runtime_checkable

import numpy as np


@runtime_checkable
class CountableSerial(Protocol):
    def get_size(self) -> int: ...


@runtime_checkable
class Mergeable(Protocol):
    def merge(self, others: Iterable[Any]) -> Any: ...


class LearningParameters:
    """Base of parameter containers: a flat fp64 view, a size in bytes, addition and scaling (what the parameter
    servers need to average or accumulate models)."""

    def flat(self) -> np.ndarray:
        raise NotImplementedError

    def get_size(self) -> int:
        return 8 * int(self.flat().size)

    def __len__(self) -> int:
        return int(self.flat().size)


class VectorBias(LearningParameters):
    """(weights: dense fp64 vector, bias: scalar) - the parameter type of the PA / SVM family."""

    def __init__(self, weights, bias: float = 0.0):
        self.weights = np.asarray(weights, dtype=np.float64).reshape(-1).copy()
        self.bias = float(bias)

    @classmethod
    def from_flat(cls, w) -> "VectorBias":
        """From the learners' flat layout `[w_0 .. w_{d-1} | b]` (bias = weight of the constant-1 feature)."""
        w = np.asarray(w, dtype=np.float64).reshape(-1)
        return cls(w[:-1], w[-1])

    @classmethod
    def from_parameters(cls, p: Dict[str, Any]) -> "VectorBias":
        """From the `parameters` map of a response / a warm-start request: {"a": [...], "b": x}."""
        return cls(np.atleast_1d(np.asarray(p["a"], dtype=np.float64)), float(p.get("b", 0.0)))

    def flat(self) -> np.ndarray:
        return np.concatenate([self.weights, [self.bias]])

    def to_parameters(self) -> Dict[str, Any]:
        return {"a": self.weights.tolist(), "b": self.bias}

    def __add__(self, o: "VectorBias") -> "VectorBias":
        return VectorBias(self.weights + o.weights, self.bias + o.bias)

    def __mul__(self, s: float) -> "VectorBias":
        return VectorBias(self.weights * s, self.bias * s)

    __rmul__ = __mul__

    def predict(self, x) -> float:
        return float(np.dot(self.weights, np.asarray(x, dtype=np.float64)) + self.bias)

    def __eq__(self, o) -> bool:
        return isinstance(o, VectorBias) and np.array_equal(self.weights, o.weights) and self.bias == o.bias

    def __repr__(self) -> str:
        return f"VectorBias(d={self.weights.size}, bias={self.bias:.6g})"


def integer_parsing(conf: Optional[Dict[str, Any]], key: str, default: int) -> int:
    """`Parsing.IntegerParsing(map, key, default)`: the value as an int, the default when absent or not a number."""
    try:
        v = (conf or {}).get(key)
        return default if v is None else int(float(v))
    except (TypeError, ValueError):
        return default


def double_parsing(conf: Optional[Dict[str, Any]], key: str, default: float) -> float:
    try:
        v = (conf or {}).get(key)
        return default if v is None else float(v)
    except (TypeError, ValueError):
        return default
