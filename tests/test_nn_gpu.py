"""The MLP learner on the GPU (tcgen05 GEMMs + the fused element-wise kernels of csrc/nn.cu) against the CPU engine on
the same stream (the glue kernels also run on the CPU emulator: tests/test_emu_kernels.py::test_nn_glue_kernels)."""
import pytest
import torch

from omldm_b200.engine import Engine
from omldm_b200.ops import functional as F
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters

pytestmark = pytest.mark.gpu


def _run(dev, hyper, kind, n_aux, impl="auto"):
    d, B = 20, 256
    with F.use_impl(impl):
        e = Engine(JobParameters(parallelism=1, testSetSize=32), device=dev)
        e.handle_request({"id": 1, "request": "Create", "learner": {"name": "NN", "hyperParameters": hyper},
                          "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
        for i in range(10):
            e.train_batch(R.generate_batch(kind, d, B, seed=9, batch_index=i, n_aux=n_aux).to(dev), d)
        l = e.pipelines[1].learner
        out = (l.buf.detach().cpu().clone(), l.counters())
        e.close()
    return out


@pytest.mark.parametrize("hyper,kind,n_aux", [({"hiddenLayers": [16, 8], "learningRate": 0.05}, R.KIND_REGRESSION, 0),
                                              ({"hiddenLayers": [16], "learningRate": 0.05, "nClasses": 4}, R.KIND_MULTICLASS, 4)])
def test_nn_gpu_matches_cpu_engine(hyper, kind, n_aux):
    """Three runs of the same stream: CPU engine (fp32), the GPU engine with plain PyTorch ops and tf32 matmuls (the
    control: how far ten SGD steps through ReLU kinks drift under tf32 rounding at all), and the native GPU path
    (tcgen05 tf32 GEMMs + csrc/nn.cu).  The native path must stay as close to the CPU run as the control does."""
    wc, cc = _run("cpu", hyper, kind, n_aux)
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        wt, ct = _run("cuda", hyper, kind, n_aux, impl="torch")
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old
    wg, cg = _run("cuda", hyper, kind, n_aux)
    assert cc["fitted"] == cg["fitted"] == ct["fitted"] > 0
    rel = float((wc - wg).norm() / wc.norm())
    ctl = float((wc - wt).norm() / wc.norm())
    assert rel < max(1e-2, 3.0 * ctl), (rel, ctl)
    assert abs(cc["loss_sum"] - cg["loss_sum"]) <= max(2e-2, 3.0 * abs(cc["loss_sum"] - ct["loss_sum"]) / abs(cc["loss_sum"])) \
        * abs(cc["loss_sum"]), (cc["loss_sum"], cg["loss_sum"], ct["loss_sum"])
