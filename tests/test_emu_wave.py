"""Kernels whose blocks meet in global memory (accumulators, "last block finishes the job" counters), re-run through
the CUDA-on-CPU emulation with the blocks of a launch running CONCURRENTLY (OMLDM_EMU_WAVE, tools/cuda_emu/README.md).
The default emulation runs blocks one after the other, which hides every inter-block ordering problem; this mode found
the missing block barrier before the completion counter of linear_minibatch_kernel / poly_scale_linear_kernel
(DESIGN.md).  The wave size is read once per process, hence the subprocess."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.join(os.path.dirname(__file__), "..")
SELECTION = ("linear_minibatch or fused_poly or kmeans or standard_scaler or sparse_mcpa or hoeffding or mcpa_coef or "
             "sq_residual or nn_glue or holdout_swap or stage_rows or f16s_row or holdout_exchange or ridge_chol")

pytestmark = pytest.mark.skipif(torch.cuda.is_available(), reason="the emulator is for machines without a GPU")


@pytest.mark.parametrize("wave", [4, 7] if os.environ.get("OMLDM_SANITIZERS_FULL") else [4])
def test_inter_block_kernels_with_concurrent_blocks(wave):
    env = dict(os.environ, OMLDM_EMU_WAVE=str(wave), OMLDM_EMU_JITTER="20", PYTHONPATH=ROOT)
    p = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_emu_kernels.py"), "-x", "-q",
                        "-p", "no:cacheprovider", "-k", SELECTION],
                       capture_output=True, text=True, env=env, cwd=ROOT, timeout=1800)
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    assert " passed" in p.stdout and "no tests ran" not in p.stdout


@pytest.mark.parametrize("with_barrier", [1, 0])
def test_control_last_block_idiom_without_the_barrier_is_caught(with_barrier):
    """The idiom of the kernels above in 15 lines (tools/cuda_emu/emu_host.cpp): with the block barrier before the
    completion counter every launch is exact, without it concurrent blocks lose additions in some launches - which
    proves that this mode sees the class of bug it is here for."""
    code = (
        "import torch\n"
        "from omldm_b200.ops import _lib\n"
        "k = _lib.use_emulator()\n"
        "bad = 0\n"
        f"for it in range({80 if with_barrier else 240}):\n"
        "    scratch, out = torch.zeros(65), torch.zeros(64)\n"
        f"    k.omldm_emu_selftest_lastblock(_lib.ptr(scratch), _lib.ptr(out), 8, {with_barrier})\n"
        "    bad += int((out != 8).any())\n"
        f"    if bad and not {with_barrier}: break\n"
        "print('bad', bad)\n")
    env = dict(os.environ, OMLDM_EMU_WAVE="8", OMLDM_EMU_JITTER="50", PYTHONPATH=ROOT)
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, cwd=ROOT, timeout=900)
    assert p.returncode == 0, p.stderr[-2000:]
    bad = int(p.stdout.split("bad")[-1])
    assert bad == 0 if with_barrier else bad > 0, p.stdout
