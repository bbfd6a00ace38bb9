"""ctypes loader for the in-tree native libraries.

`libomldm_kernels.so` exposes an `extern "C"` ABI (declared in `include/omldm_kernels.h`) that
takes raw device pointers and a `cudaStream_t`; Python hands it `tensor.data_ptr()` and the
current torch stream.  There is deliberately no silent fallback on a GPU machine: if CUDA is
available and the library cannot be loaded, every op raises.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

import torch

_LIBDIR = Path(__file__).resolve().parents[1] / "lib"
_kernels = None
_host = None
_emulated = False


class NativeError(RuntimeError):
    pass


def _try_build() -> None:
    if os.environ.get("OMLDM_NO_AUTOBUILD"):
        return
    try:
        from . import build

        build.build_all()
    except Exception as e:  # pragma: no cover - reported by the caller
        raise NativeError(f"building the native libraries failed: {e}") from e


def kernels() -> ctypes.CDLL:
    """The CUDA kernel library (requires a CUDA runtime to be loadable)."""
    global _kernels
    if _kernels is None:
        path = _LIBDIR / "libomldm_kernels.so"
        if os.environ.get("OMLDM_KERNELS_LIB"):          # kernel-tuning experiments: an alternative build of the library
            path = Path(os.environ["OMLDM_KERNELS_LIB"])
        if not path.exists():
            _try_build()
        if not path.exists():
            raise NativeError(f"{path} is missing; run `python -m omldm_b200.ops.build`")
        try:
            _kernels = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
        except OSError as e:
            raise NativeError(f"cannot load {path}: {e}") from e
        _kernels.omldm_last_error.restype = ctypes.c_char_p
        _kernels.omldm_launch_count.restype = ctypes.c_longlong
    return _kernels


def use_emulator(asan: bool = False, tsan: bool = False) -> ctypes.CDLL:
    """Route `kernels()` to the CUDA-on-CPU emulation of the PTX-free kernel files (tools/cuda_emu): the real
    kernel sources compiled by g++, one host thread per CUDA thread, operating on host tensors.  For tests on
    machines without a GPU (`functional.use_impl("emu")`); never used on a GPU box."""
    global _kernels, _emulated
    import importlib.util
    root = Path(__file__).resolve().parents[2]
    spec = importlib.util.spec_from_file_location("omldm_build_emu", root / "tools" / "cuda_emu" / "build_emu.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    tsan = tsan or bool(os.environ.get("OMLDM_EMU_TSAN"))
    out = mod.OUT / ("libomldm_emu_tsan.so" if tsan else "libomldm_emu_asan.so" if asan else "libomldm_emu.so")
    srcs = [root / "csrc" / s for s in mod.SOURCES] + list((root / "tools" / "cuda_emu").glob("*"))
    if not out.exists() or any(p.stat().st_mtime > out.stat().st_mtime for p in srcs):
        mod.build(asan, tsan)
    _kernels = ctypes.CDLL(str(out))
    _kernels.omldm_last_error.restype = ctypes.c_char_p
    _kernels.omldm_launch_count.restype = ctypes.c_longlong
    _emulated = True
    return _kernels


def stop_emulator() -> None:
    global _kernels, _emulated
    if _emulated:
        _kernels, _emulated = None, False


def host() -> ctypes.CDLL:
    """The host runtime library (ring buffers, record parser); no CUDA dependency."""
    global _host
    if _host is None:
        path = _LIBDIR / "libomldm_host.so"
        if not path.exists():
            _try_build()
        if not path.exists():
            raise NativeError(f"{path} is missing; run `python -m omldm_b200.ops.build`")
        _host = ctypes.CDLL(str(path))
    return _host


def have_cuda() -> bool:
    return torch.cuda.is_available()


def use_native(*tensors: torch.Tensor) -> bool:
    """True when the tensors live on a GPU (then the native path is mandatory)."""
    return any(t is not None and t.is_cuda for t in tensors)


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = kernels().omldm_last_error().decode(errors="replace")
        raise NativeError(f"{what} failed: {msg}")


def stream_ptr() -> ctypes.c_void_p:
    if _emulated:
        return ctypes.c_void_p(0)
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t: torch.Tensor | None) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


_replayed = 0        # kernels launched through graph replays, kept on the Python side (no native call per step)


def launch_count() -> int:
    """Number of kernels this library has launched since the last reset (0 on CPU-only hosts): direct launches
    counted by the library plus the kernels of replayed CUDA graphs."""
    if not have_cuda():
        return 0
    return int(kernels().omldm_launch_count()) + _replayed


def add_launch_count(n: int) -> None:
    """Account for kernels launched by replaying a captured CUDA graph (n per replay; negative to take back the
    launches counted while capturing)."""
    global _replayed
    _replayed += n


def reset_launch_count() -> None:
    global _replayed
    _replayed = 0
    if have_cuda():
        kernels().omldm_reset_launch_count()
