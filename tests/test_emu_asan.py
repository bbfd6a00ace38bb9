"""Memory-safety check of the kernel sources: tests/test_emu_kernels.py and tests/test_emu_sync.py re-run in a subprocess with the
AddressSanitizer + UBSan build of the CUDA-on-CPU emulation (tools/cuda_emu).  `compute-sanitizer` is not available on
the GPU pool this project is developed on, so this is where out-of-bounds accesses to global, shared and
dynamic shared memory of the PTX-free kernels are caught.  A negative control proves the sanitizer is live."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.join(os.path.dirname(__file__), "..")


def _asan_env():
    if os.environ.get("OMLDM_SKIP_EMU_ASAN"):
        pytest.skip("OMLDM_SKIP_EMU_ASAN is set (the sanitizer run takes about four minutes)")
    lib = subprocess.run(["gcc", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
    if not lib or not os.path.isabs(lib) or not os.path.exists(lib):
        pytest.skip("libasan is not installed")
    env = dict(os.environ, OMLDM_EMU_ASAN="1", LD_PRELOAD=lib, ASAN_OPTIONS="detect_leaks=0:halt_on_error=1",
               UBSAN_OPTIONS="print_stacktrace=1:halt_on_error=1", PYTHONPATH=ROOT)
    return env


@pytest.mark.skipif(torch.cuda.is_available(), reason="the emulator is for machines without a GPU")
def test_kernels_are_clean_under_asan_ubsan():
    env = _asan_env()
    p = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_emu_kernels.py"),
                        os.path.join(ROOT, "tests", "test_emu_sync.py"), "-x", "-q", "-p", "no:cacheprovider",
                        # engine-level tests re-run the same kernels; the 2-blocks-per-rank FGM case is the slow one
                        "-k", "not engine_pipelines and not (fgm_rounds and 2)"],
                       capture_output=True, text=True, env=env, cwd=ROOT, timeout=2400)
    tail = (p.stdout + p.stderr)[-3000:]
    assert p.returncode == 0, tail
    assert "AddressSanitizer" not in tail and "runtime error" not in tail, tail


@pytest.mark.skipif(torch.cuda.is_available(), reason="the emulator is for machines without a GPU")
def test_negative_control_overrun_is_caught():
    env = _asan_env()
    code = (
        "import torch, ctypes\n"
        "from omldm_b200.ops import _lib\n"
        "k = _lib.use_emulator(asan=True)\n"
        "raw = torch.randn(8, 11)\n"
        "out = torch.empty(8 * 32 - 20)\n"          # 20 floats short of 8 staged rows of pitch 32
        "k.omldm_stage_rows(_lib.ptr(raw), 11, 8, 10, _lib.ptr(out), 32, None)\n"
        "print('survived')\n")
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)
    assert p.returncode != 0 and "AddressSanitizer" in p.stderr and "survived" not in p.stdout, p.stderr[-2000:]
