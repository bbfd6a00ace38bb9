"""The MLP learner on the GPU (tcgen05 GEMMs + the fused element-wise kernels of csrc/nn.cu) against the CPU engine on
the same stream (the glue kernels also run on the CPU emulator: tests/test_emu_kernels.py::test_nn_glue_kernels)."""
import pytest
import torch

from omldm_b200.engine import Engine
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("hyper,kind,n_aux", [({"hiddenLayers": [16, 8], "learningRate": 0.05}, R.KIND_REGRESSION, 0),
                                              ({"hiddenLayers": [16], "learningRate": 0.05, "nClasses": 4}, R.KIND_MULTICLASS, 4)])
def test_nn_gpu_matches_cpu_engine(hyper, kind, n_aux):
    d, B = 20, 256
    out = []
    for dev in ("cpu", "cuda"):
        e = Engine(JobParameters(parallelism=1, testSetSize=32), device=dev)
        e.handle_request({"id": 1, "request": "Create", "learner": {"name": "NN", "hyperParameters": hyper},
                          "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
        for i in range(10):
            e.train_batch(R.generate_batch(kind, d, B, seed=9, batch_index=i, n_aux=n_aux).to(dev), d)
        l = e.pipelines[1].learner
        out.append((l.buf.detach().cpu().clone(), l.counters()))
        e.close()
    (wc, cc), (wg, cg) = out
    assert cc["fitted"] == cg["fitted"] > 0
    # tf32 products on the GPU, carried through 10 SGD steps and ReLU masks: a unit that sits at its kink may fall on
    # the other side, so single weights can move by more than their rounding error; the model as a whole must agree
    rel = float((wc - wg).norm() / wc.norm())
    worst = float((wc - wg).abs().max() / wc.abs().max())
    assert rel < 1e-2 and worst < 5e-2, (rel, worst)
    assert abs(cc["loss_sum"] - cg["loss_sum"]) <= 2e-2 * abs(cc["loss_sum"])
