"""Device preprocessors on staged rows (whitelist PipelineMap.scala:67; kernels csrc/preprocess.cu)."""
from __future__ import annotations

from typing import Any, Dict, Optional

import torch

from ..ops import functional as F
from ..oracle.learners import _h
from ..oracle.preprocessors import poly_factor_table


class GpuPreprocessor:
    name = ""

    def __init__(self, hyper: Optional[dict], device):
        self.hyper = dict(hyper or {})
        self.device = torch.device(device)
        self.d_in: Optional[int] = None

    def out_dim(self, d: int) -> int:
        return d

    def bind(self, d: int) -> None:
        self.d_in = d

    def apply(self, Z: torch.Tensor, fit: bool) -> torch.Tensor:
        """Transforms the batch (in place when the width is unchanged); returns the batch to use."""
        raise NotImplementedError

    def to_parameters(self) -> Dict[str, Any]:
        return {}

    def state_dict(self) -> dict:
        return {}

    def load_state_dict(self, sd: dict) -> None:
        pass


class GpuPolynomialFeatures(GpuPreprocessor):
    name = "PolynomialFeatures"

    def __init__(self, hyper, device):
        super().__init__(hyper, device)
        self.degree = int(_h(hyper, ["degree"], 2))

    def out_dim(self, d):
        return len(poly_factor_table(d, self.degree))

    def bind(self, d):
        super().bind(d)
        tab = poly_factor_table(d, self.degree)
        self.d_out = len(tab)
        self.factors = torch.from_numpy(tab).to(self.device)

    def apply(self, Z, fit):
        return F.poly_features(Z, self.d_in, self.factors, self.d_out)


class GpuStandardScaler(GpuPreprocessor):
    name = "StandardScaler"

    def bind(self, d):
        super().bind(d)
        self.state = torch.zeros(1 + 2 * d, dtype=torch.float64, device=self.device)
        self.scratch = torch.zeros(3 + 4 * d, dtype=torch.float64, device=self.device)

    def apply(self, Z, fit):
        F.standard_scaler(Z, self.d_in, self.state, self.scratch, fit)
        return Z

    def to_parameters(self):
        s = self.state.cpu().numpy()
        d = self.d_in
        return {"count": float(s[0]), "mean": s[1:1 + d].tolist(), "m2": s[1 + d:1 + 2 * d].tolist()}

    def state_dict(self):
        return {"state": self.state.cpu().clone()}

    def load_state_dict(self, sd):
        self.state.copy_(sd["state"].to(self.device))


class GpuMinMaxScaler(GpuPreprocessor):
    name = "MinMaxScaler"

    def bind(self, d):
        super().bind(d)
        self.state = torch.empty(2 * d, dtype=torch.float32, device=self.device)
        self.state[:d] = float("inf")
        self.state[d:] = float("-inf")

    def apply(self, Z, fit):
        F.minmax_scaler(Z, self.d_in, self.state, fit)
        return Z

    def to_parameters(self):
        s = self.state.cpu().numpy()
        d = self.d_in
        return {"min": s[:d].tolist(), "max": s[d:].tolist()}

    def state_dict(self):
        return {"state": self.state.cpu().clone()}

    def load_state_dict(self, sd):
        self.state.copy_(sd["state"].to(self.device))


GPU_PREPROCESSORS = {c.name: c for c in (GpuPolynomialFeatures, GpuStandardScaler, GpuMinMaxScaler)}


def make_gpu_preprocessor(name: str, hyper, device) -> GpuPreprocessor:
    return GPU_PREPROCESSORS[name](hyper, device)
