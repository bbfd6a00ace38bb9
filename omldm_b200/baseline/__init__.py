"""The in-repo PyTorch + NCCL (+ cuBLAS) baseline: the number the fused sm_100a path has to beat.

The reference publishes no measurements and cannot be built offline, so `BASELINE.json`'s fallback applies:
"its NCCL(+cuBLAS) equivalent ... a path that only calls NCCL is the baseline, not the product".  The baseline is the
SAME engine with two switches flipped:

* `ops.functional.set_impl("torch")` - every op (holdout, preprocessors, learners, GEMMs through cuBLAS tf32) runs as
  stock PyTorch instead of the hand-written kernels;
* `Engine(..., impl="nccl")` - every protocol step is `torch.distributed` collectives plus torch glue instead of the
  fused peer-memory kernels (`parallel/fused.py`, the non-fused branch of each driver).

`bench.py --impl nccl` measures it; `tests/test_multigpu.py` checks that both produce the same models.
"""
from __future__ import annotations

import contextlib


@contextlib.contextmanager
def torch_ops():
    """Run the enclosed code with the stock-PyTorch implementation of every op."""
    from ..ops import functional as F
    with F.use_impl("torch"):
        yield


def make_engine(params=None, rank: int = 0, world: int = 1, device=None, group=None):
    """An `Engine` whose protocol steps use torch.distributed collectives (NCCL on GPUs, gloo on CPU).  Combine with
    `torch_ops()` for the complete baseline."""
    from ..engine import Engine
    return Engine(params, rank, world, device, impl="nccl", group=group)
