"""Hoeffding tree grown on the device (csrc/tree.cu::omldm_ht_split) against the host-grown tree of the CPU engine.
First hardware run pending (written after round 1's GPU budget was spent); the kernels have run on the CPU emulator
(tests/test_emu_kernels.py::test_ht_split_kernel_matches_the_host_split_test)."""
import pytest

from omldm_b200.engine import Engine
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters

pytestmark = pytest.mark.gpu


def test_device_grown_tree_matches_host_grown_tree():
    d, B, C = 6, 512, 3
    hyper = {"nClasses": C, "gracePeriod": 100, "delta": 1e-3, "tieThreshold": 0.2}
    out = []
    for dev in ("cpu", "cuda"):
        e = Engine(JobParameters(parallelism=1, testSetSize=16), device=dev)
        e.handle_request({"id": 1, "request": "Create", "learner": {"name": "HT", "hyperParameters": hyper},
                          "trainingConfiguration": {"protocol": "Synchronous", "nFeatures": d}})
        for i in range(8):
            e.train_batch(R.generate_batch(R.KIND_CLUSTERS, d, B, seed=31, batch_index=i, n_aux=C).to(dev), d)
        l = e.pipelines[1].learner
        t = l.to_oracle()
        probe = R.generate_batch(R.KIND_CLUSTERS, d, 256, seed=31, batch_index=20, n_aux=C).to(dev)
        out.append((list(t.feature), list(t.threshold), l.counters()["fitted"], l.predict(probe).cpu()))
        e.close()
    (fc, tc, nc, pc), (fg, tg, ng, pg) = out
    assert nc == ng and len(fc) >= 3 and len(fg) >= 3
    assert fc[0] == fg[0] and abs(tc[0] - tg[0]) < 1e-3 * max(1.0, abs(tc[0]))
    assert float((pc == pg).float().mean()) >= 0.9
