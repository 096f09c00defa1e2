"""In-tree native build for idc_models_b200.

Two shared libraries are produced under ``idc_models_b200/_native/``:

* ``libidcb200.so``      – every CUDA kernel (``csrc/*.cu``), compiled by nvcc for
  ``-gencode arch=compute_100a,code=sm_100a`` (tcgen05/TMEM/TMA need the ``a`` target) with
  ``-lineinfo`` so ncu source pages map back to the .cu files. cudart is linked statically, so the
  library has no dependency beyond the CUDA driver.
* ``libidcb200_host.so`` – host-only C++ (``csrc/host/*.cpp``): the minimal HDF5 reader/writer,
  the flat checkpoint writer and the pinned prefetch ring. It has no CUDA dependency and is
  loadable on a CPU-only box (the unit tests rely on that).

nvcc cross-compiles without a GPU, so ``build()`` is also the "does it compile" check the driver
runs on the CPU box. Objects are cached by source mtime under ``build/``.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
OUT = ROOT / "idc_models_b200" / "_native"
OBJ = ROOT / "build"
CUDA_LIB = OUT / "libidcb200.so"
HOST_LIB = OUT / "libidcb200_host.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=default",
]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-pthread"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; set NVCC or install the CUDA toolkit")


def _newer(src: Path, dst: Path, extra: list[Path]) -> bool:
    if not dst.exists():
        return True
    t = dst.stat().st_mtime
    return any(p.stat().st_mtime > t for p in [src, *extra])


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError(f"native build failed: {cmd[0]} {cmd[-1]}")


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    nvcc = _nvcc()
    OBJ.mkdir(exist_ok=True)
    OUT.mkdir(exist_ok=True)
    headers = list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    srcs = sorted(CSRC.glob("*.cu"))
    jobs = []
    objs = []
    for s in srcs:
        o = OBJ / (s.stem + ".o")
        objs.append(o)
        if force or _newer(s, o, headers):
            cmd = [nvcc, *NVCC_FLAGS, "-c", str(s), "-o", str(o)]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(_run, jobs))
    if jobs or not CUDA_LIB.exists() or any(o.stat().st_mtime > CUDA_LIB.stat().st_mtime for o in objs):
        _run([nvcc, "-shared", "-o", str(CUDA_LIB), *map(str, objs), "-cudart", "static",
              "-gencode", "arch=compute_100a,code=sm_100a"])
    return CUDA_LIB


def build_host(force: bool = False) -> Path:
    cxx = shutil.which("g++") or "g++"
    OBJ.mkdir(exist_ok=True)
    OUT.mkdir(exist_ok=True)
    srcs = sorted((CSRC / "host").glob("*.cpp"))
    headers = list((CSRC / "host").glob("*.h")) + list((ROOT / "include").glob("*.h"))
    objs = []
    jobs = []
    for s in srcs:
        o = OBJ / ("host_" + s.stem + ".o")
        objs.append(o)
        if force or _newer(s, o, headers):
            jobs.append([cxx, *CXX_FLAGS, "-I", str(ROOT / "include"), "-c", str(s), "-o", str(o)])
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(_run, jobs))
    if (jobs or not HOST_LIB.exists()) and objs:
        _run([cxx, "-shared", "-pthread", "-o", str(HOST_LIB), *map(str, objs)])
    return HOST_LIB


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_host(force)
    build_cuda(force, verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print("built", CUDA_LIB, HOST_LIB)
