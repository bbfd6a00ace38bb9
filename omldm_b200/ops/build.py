"""In-tree native build: `python -m omldm_b200.ops.build`.

Compiles every `csrc/*.cu` for sm_100a (and only sm_100a) into
`omldm_b200/lib/libomldm_kernels.so`, and `csrc/host/*.cpp` into
`omldm_b200/lib/libomldm_host.so`.  nvcc cross-compiles without a GPU, so this runs on the
CPU build box; the resulting `.so` files travel to the GPU box with the source tree.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
LIBDIR = ROOT / "omldm_b200" / "lib"
OBJDIR = ROOT / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-I", str(ROOT / "include"), "-I", str(CSRC),
]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-pthread", "-I", str(ROOT / "include")]


def _nvcc() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found; the CUDA toolkit is required to build omldm_b200")
    return cand


def _stale(src: Path, out: Path, extra_deps: list[Path]) -> bool:
    if not out.exists():
        return True
    t = out.stat().st_mtime
    return any(p.stat().st_mtime > t for p in [src, *extra_deps])


def _run(cmd: list[str]) -> str:
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError("command failed: %s\n%s\n%s" % (" ".join(cmd), p.stdout, p.stderr))
    return p.stdout + p.stderr


def build_kernels(verbose: bool = False, ptxas_info: bool = False) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    OBJDIR.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()
    headers = list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list((ROOT / "include").glob("*.h"))
    sources = sorted(CSRC.glob("*.cu"))
    out = LIBDIR / "libomldm_kernels.so"
    jobs = []
    for src in sources:
        obj = OBJDIR / (src.stem + ".o")
        if _stale(src, obj, headers):
            cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
            if ptxas_info:
                cmd += ["-Xptxas", "-v"]
            jobs.append(cmd)
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for log in ex.map(_run, jobs):
                if verbose or ptxas_info:
                    print(log)
    objs = [OBJDIR / (s.stem + ".o") for s in sources]
    if jobs or not out.exists():
        _run([nvcc, "-shared", "-cudart", "shared", "-o", str(out), *map(str, objs)])
    return out


def build_host(verbose: bool = False) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    OBJDIR.mkdir(parents=True, exist_ok=True)
    out = LIBDIR / "libomldm_host.so"
    sources = sorted(s for s in (CSRC / "host").glob("*.cpp") if not s.name.endswith("_test.cpp"))
    headers = list((ROOT / "include").glob("*.h")) + list((CSRC / "host").glob("*.h"))
    if not sources:
        return out
    if any(_stale(s, out, headers) for s in sources):
        log = _run(["g++", *CXX_FLAGS, "-shared", "-o", str(out), *map(str, sources)])
        if verbose:
            print(log)
    return out


def build_all(verbose: bool = False, ptxas_info: bool = False) -> None:
    build_host(verbose)
    build_kernels(verbose, ptxas_info)


if __name__ == "__main__":
    build_all(verbose=True, ptxas_info="--ptxas" in sys.argv)
    print("built:", *(p.name for p in LIBDIR.glob("*.so")))
