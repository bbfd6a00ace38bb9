"""Sanitizer runs of the kernel sources through the CUDA-on-CPU emulation (tools/cuda_emu): tests/test_emu_kernels.py
and tests/test_emu_sync.py re-run in subprocesses against the AddressSanitizer + UBSan build (out-of-bounds accesses to
global, shared and dynamic shared memory) and the ThreadSanitizer build (data races: a missing __syncthreads, a flag
protocol whose release / acquire pairing does not order the data it guards).  `compute-sanitizer` is not available on
the GPU pool this project is developed on; this is its stand-in.  Negative controls prove both sanitizers
are live.

By default a quick selection runs (the kernels written last, the core of the protocol kernels); set
OMLDM_SANITIZERS_FULL=1 for everything, including ThreadSanitizer with concurrent blocks (about fifteen minutes; the log of such a run is profiles/sanitizers_emu.txt) or
OMLDM_SKIP_EMU_SANITIZERS=1 to skip."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.join(os.path.dirname(__file__), "..")
FULL = bool(os.environ.get("OMLDM_SANITIZERS_FULL"))
QUICK = ("stage_rows_f16 or sparse_mcpa or ridge_cg or holdout_swap or edge_shapes or ht_split or sum_delta or gm_votes or "
         "async_and_easgd or barrier or forwarding or scatter_push or nn_glue")
FULL_SEL = "not engine_pipelines and not (fgm_rounds and 2)"

pytestmark = pytest.mark.skipif(torch.cuda.is_available(), reason="the emulator is for machines without a GPU")


def _env(kind):
    if os.environ.get("OMLDM_SKIP_EMU_SANITIZERS") or os.environ.get("OMLDM_SKIP_EMU_ASAN"):
        pytest.skip("sanitizer runs skipped by the environment")
    name = "libasan.so" if kind == "asan" else "libtsan.so"
    lib = subprocess.run(["gcc", "-print-file-name=" + name], capture_output=True, text=True).stdout.strip()
    if not lib or not os.path.isabs(lib) or not os.path.exists(lib):
        pytest.skip(name + " is not installed")
    env = dict(os.environ, LD_PRELOAD=lib, PYTHONPATH=ROOT)
    if kind == "asan":
        env.update(OMLDM_EMU_ASAN="1", ASAN_OPTIONS="detect_leaks=0:halt_on_error=1",
                   UBSAN_OPTIONS="print_stacktrace=1:halt_on_error=1")
    else:
        # libtorch's own CPU ops run on libgomp threads, which TSan cannot see into (it reports their TensorIterator
        # bookkeeping as races); one OpenMP thread keeps the reports to the code under test: the emulated kernels
        env.update(OMLDM_EMU_TSAN="1", TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0 exitcode=66",
                   OMP_NUM_THREADS="1", MKL_NUM_THREADS="1")
    return env


def _run_suites(env):
    return subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_emu_kernels.py"),
                           os.path.join(ROOT, "tests", "test_emu_sync.py"), "-x", "-q", "-p", "no:cacheprovider",
                           "-k", FULL_SEL if FULL else QUICK],
                          capture_output=True, text=True, env=env, cwd=ROOT, timeout=3000)


def test_kernels_are_clean_under_asan_ubsan():
    p = _run_suites(_env("asan"))
    tail = (p.stdout + p.stderr)[-3000:]
    assert p.returncode == 0, tail
    assert "AddressSanitizer" not in tail and "runtime error" not in tail, tail


def test_negative_control_overrun_is_caught():
    code = (
        "import torch, ctypes\n"
        "from omldm_b200.ops import _lib\n"
        "k = _lib.use_emulator(asan=True)\n"
        "raw = torch.randn(8, 11)\n"
        "out = torch.empty(8 * 32 - 20)\n"          # 20 floats short of 8 staged rows of pitch 32
        "k.omldm_stage_rows(_lib.ptr(raw), 11, 8, 10, _lib.ptr(out), 32, None)\n"
        "print('survived')\n")
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=_env("asan"), cwd=ROOT,
                       timeout=900)
    assert p.returncode != 0 and "AddressSanitizer" in p.stderr and "survived" not in p.stdout, p.stderr[-2000:]


def test_kernels_are_race_free_under_tsan():
    p = _run_suites(_env("tsan"))
    out = p.stdout + p.stderr
    assert "WARNING: ThreadSanitizer" not in out, out[-4000:]
    assert p.returncode == 0, out[-3000:]


@pytest.mark.parametrize("with_barrier", [1, 0])
def test_negative_control_missing_syncthreads_is_caught(with_barrier):
    code = (
        "import torch, sys\n"
        "from omldm_b200.ops import _lib\n"
        "k = _lib.use_emulator(tsan=True)\n"
        "out = torch.zeros(64)\n"
        f"k.omldm_emu_selftest_shared(_lib.ptr(out), {with_barrier})\n"
        "print('ran', out[:3].tolist())\n")
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=_env("tsan"), cwd=ROOT,
                       timeout=900)
    assert "ran" in p.stdout
    if with_barrier:
        assert p.returncode == 0 and "ThreadSanitizer" not in p.stderr, p.stderr[-2000:]
        assert "[1.0, 2.0, 3.0]" in p.stdout
    else:
        assert p.returncode == 66 and "data race" in p.stderr, p.stderr[-2000:]


@pytest.mark.parametrize("with_barrier", [1, 0])
def test_negative_control_unordered_last_block_is_caught(with_barrier):
    """Inter-block ordering: blocks run concurrently (OMLDM_EMU_WAVE) and the floating-point atomics of the shim are
    relaxed like CUDA's, so a last block that reads / clears an accumulator without being ordered after every adder
    is a ThreadSanitizer report in every run - the bug fixed in linear_minibatch_kernel / poly_scale_linear_kernel."""
    code = (
        "import torch\n"
        "from omldm_b200.ops import _lib\n"
        "k = _lib.use_emulator(tsan=True)\n"
        "for it in range(10):\n"
        "    scratch, out = torch.zeros(65), torch.zeros(64)\n"
        f"    k.omldm_emu_selftest_lastblock(_lib.ptr(scratch), _lib.ptr(out), 8, {with_barrier})\n"
        "print('ran')\n")
    env = dict(_env("tsan"), OMLDM_EMU_WAVE="8")
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, cwd=ROOT, timeout=900)
    assert "ran" in p.stdout
    if with_barrier:
        assert p.returncode == 0 and "ThreadSanitizer" not in p.stderr, p.stderr[-2000:]
    else:
        assert p.returncode == 66 and "data race" in p.stderr, p.stderr[-2000:]


@pytest.mark.skipif(not FULL, reason="set OMLDM_SANITIZERS_FULL=1 (about four minutes)")
def test_kernels_are_race_free_under_tsan_with_concurrent_blocks():
    """Everything again under ThreadSanitizer with the blocks of plain launches running concurrently
    (OMLDM_EMU_WAVE=4): inter-block ordering, and the emulated ranks of the asynchronous protocol kernels at once."""
    p = _run_suites(dict(_env("tsan"), OMLDM_EMU_WAVE="4"))
    out = p.stdout + p.stderr
    assert "WARNING: ThreadSanitizer" not in out, out[-4000:]
    assert p.returncode == 0, out[-3000:]
