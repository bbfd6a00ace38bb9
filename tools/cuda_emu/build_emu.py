"""Builds build/emu/libomldm_emu.so: the PTX-free kernel files of csrc/ compiled by g++ against the CUDA-on-CPU
shim (emu_cuda.h), exposing the same `omldm_*` C ABI on host pointers.

    python tools/cuda_emu/build_emu.py            # plain build, used by tests/test_emu_kernels.py
    python tools/cuda_emu/build_emu.py --asan     # AddressSanitizer + UBSan build (libomldm_emu_asan.so)

Source rewriting (nothing else is touched):
  kernel<<<grid, block, smem, stream>>>(args)  ->  emu::launcher(kernel, grid, block, smem, stream)(args)
  extern __shared__ T name[];                  ->  T* name = reinterpret_cast<T*>(emu::dyn_smem());
  __shared__ T a[N], b;                        ->  T (&a)[N] = *emu::block_shared<T[N]>(id); T& b = ...  (per block)
Headers <cuda.h>, <cuda_runtime.h>, <cuda_fp16.h>, <cooperative_groups.h>, <math_constants.h>, "ptx.cuh" and
"host_common.h" resolve to stand-ins in the build directory.
"""
from __future__ import annotations

import re
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
HERE = Path(__file__).resolve().parent
OUT = ROOT / "build" / "emu"
# kernel files whose PTX lives behind ptx.cuh helpers that have a host meaning (sys-scope loads / stores / atomics
# become host atomics); umma_gemm.cu (tcgen05 / TMA) is hardware only
SOURCES = ["ingest.cu", "sparse_mcpa.cu", "cluster.cu", "linear_learners.cu", "preprocess.cu", "tree.cu",
           "fused_pipeline.cu", "ridge_cg.cu", "ridge_chol.cu", "rows_f16.cu", "sync.cu", "nn.cu"]

_LAUNCH = re.compile(r"([A-Za-z_][\w:]*(?:<[^<>;(){}]*>)?)\s*<<<(.+?)>>>", re.S)
_DYN_SMEM = re.compile(r"extern\s+__shared__\s+([\w:]+)\s+(\w+)\s*\[\s*\]\s*;")


_SHARED = re.compile(r"(?<!extern )__shared__\s+([\w:]+)\s+([^;]+);")
_DECL = re.compile(r"\s*(\w+)\s*((?:\[[^\]]*\])*)\s*$")


def _shared(m: "re.Match") -> str:
    """`__shared__ T a[N], b;` -> references bound to per-block storage (emu::block_shared)."""
    ty, out = m.group(1), []
    for decl in m.group(2).split(","):
        d = _DECL.match(decl)
        if d is None:
            raise RuntimeError("cannot rewrite shared declaration: " + m.group(0))
        name, dims = d.group(1), d.group(2)
        if dims:
            out.append(f"{ty} (&{name}){dims} = *emu::block_shared<{ty}{dims}>(__COUNTER__);")
        else:
            out.append(f"{ty}& {name} = *emu::block_shared<{ty}>(__COUNTER__);")
    return " ".join(out)


def rewrite(text: str) -> str:
    text = _DYN_SMEM.sub(r"\1* \2 = reinterpret_cast<\1*>(emu::dyn_smem());", text)
    text = _SHARED.sub(_shared, text)
    return _LAUNCH.sub(r"emu::launcher(\1, \2)", text)


def build(asan: bool = False, tsan: bool = False) -> Path:
    src_dir = OUT / ("src_tsan" if tsan else "src_asan" if asan else "src")
    if src_dir.exists():
        shutil.rmtree(src_dir)
    src_dir.mkdir(parents=True)
    shutil.copy(HERE / "emu_cuda.h", src_dir / "emu_cuda.h")
    shutil.copy(HERE / "stub_ptx.cuh", src_dir / "ptx.cuh")
    shutil.copy(HERE / "stub_host_common.h", src_dir / "host_common.h")
    shutil.copy(HERE / "emu_host.cpp", src_dir / "emu_host.cpp")
    for h in ("cuda.h", "cuda_runtime.h", "cuda_fp16.h", "cooperative_groups.h", "math_constants.h"):
        (src_dir / h).write_text('#pragma once\n#include "emu_cuda.h"\n')
    for h in (ROOT / "csrc").glob("*.cuh"):
        if h.name != "ptx.cuh":
            (src_dir / h.name).write_text(rewrite(h.read_text()))
    cpps = [src_dir / "emu_host.cpp"]
    for name in SOURCES:
        dst = src_dir / (Path(name).stem + ".cpp")
        dst.write_text(rewrite((ROOT / "csrc" / name).read_text()))
        cpps.append(dst)
    out = OUT / ("libomldm_emu_tsan.so" if tsan else "libomldm_emu_asan.so" if asan else "libomldm_emu.so")
    # -Bsymbolic: the real kernel library may already be loaded RTLD_GLOBAL in the same process and defines the
    # same symbols (check_launch, the kernels themselves); the emulator must call its own
    flags = ["-std=c++17", "-fPIC", "-shared", "-pthread", "-Wl,-Bsymbolic", "-I", str(src_dir), "-I", str(ROOT / "include"),
             "-Wall", "-Wno-unknown-pragmas", "-Wno-unused-variable", "-Wno-sign-compare"]
    if tsan:
        flags += ["-O1", "-g", "-fsanitize=thread", "-fno-omit-frame-pointer"]
    else:
        flags += ["-O1", "-g", "-fsanitize=address,undefined", "-fno-omit-frame-pointer"] if asan else ["-O2"]
    # one object per source, compiled in parallel, then one link
    import concurrent.futures as cf
    cflags = [f for f in flags if f not in ("-shared", "-Wl,-Bsymbolic")]

    def compile_one(src: Path):
        obj = src.with_suffix(".o")
        p = subprocess.run(["g++", *cflags, "-c", str(src), "-o", str(obj)], capture_output=True, text=True)
        return obj, p

    errs, objs = [], []
    with cf.ThreadPoolExecutor(max_workers=min(8, len(cpps))) as ex:
        for obj, p in ex.map(compile_one, cpps):
            objs.append(obj)
            if p.returncode != 0:
                errs.append(p.stdout + p.stderr)
    if errs:
        raise RuntimeError("emulator build failed:\n" + "\n".join(errs))
    p = subprocess.run(["g++", *flags, *map(str, objs), "-o", str(out)], capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError("emulator link failed:\n" + p.stdout + p.stderr)
    return out


if __name__ == "__main__":
    print(build("--asan" in sys.argv, "--tsan" in sys.argv))
