"""In-tree build of the native libraries.

* ``libdistlearn_b200.so`` — CUDA kernels + symmetric-heap plumbing, compiled by nvcc for
  sm_100a only (``-gencode arch=compute_100a,code=sm_100a -lineinfo``).  Cross-compiles on a
  machine without a GPU.
* ``libdistlearn_host.so`` — the CPU (POSIX shared memory) transport, compiled by g++; it has
  no CUDA dependency so it loads on GPU-less machines.

The libraries land in ``torch_distlearn_b200/lib/`` (git-ignored, shipped to the GPU box by
gpurun).  ``python -m torch_distlearn_b200.build`` rebuilds what is out of date.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
LIBDIR = Path(__file__).resolve().parent / "lib"
OBJDIR = ROOT / "build" / "obj"

CUDA_SOURCES = [
    "dl_host.cu",
    "heap.cu",
    "allreduce.cu",
    "fused_sgd.cu",
    "fused_ea.cu",
    "broadcast.cu",
    "async_ea.cu",
    "local_ops.cu",
    "ll_ops.cu",
    "litmus.cu",
    "peak.cu",
    "mnist_step.cu",
    "tc_gemm.cu",
    "conv_ops.cu",
    "bn_ops.cu",
    "head_ops.cu",
]
HOST_SOURCES = ["host_transport.cpp", "loader.cpp"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def _nvcc() -> str | None:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    return cand if os.path.exists(cand) else None


def _stamp(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(str(p).encode())
        h.update(Path(p).read_bytes())
    return h.hexdigest()


def _deps() -> list[Path]:
    return sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h"))
                  + list(CSRC.glob("*.cpp")) + list((ROOT / "include").glob("*.h")))


def _run(cmd: list[str], log: Path | None = None) -> None:
    res = subprocess.run(cmd, capture_output=True, text=True)
    if log is not None:
        log.parent.mkdir(parents=True, exist_ok=True)
        log.write_text(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("build step failed: " + " ".join(cmd))


def cuda_lib_path() -> Path:
    return LIBDIR / "libdistlearn_b200.so"


def host_lib_path() -> Path:
    return LIBDIR / "libdistlearn_host.so"


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    nvcc = _nvcc()
    if nvcc is None:
        raise RuntimeError("nvcc not found: cannot build libdistlearn_b200.so")
    LIBDIR.mkdir(parents=True, exist_ok=True)
    OBJDIR.mkdir(parents=True, exist_ok=True)
    out = cuda_lib_path()
    stamp_file = LIBDIR / ".cuda.stamp"
    stamp = _stamp([p for p in _deps() if p.suffix != ".cpp"])
    if out.exists() and not force and stamp_file.exists() and stamp_file.read_text() == stamp:
        return out

    def compile_one(src: str) -> Path:
        obj = OBJDIR / (src + ".o")
        _run([nvcc, *NVCC_FLAGS, "-c", str(CSRC / src), "-o", str(obj)], log=OBJDIR / (src + ".ptxas.log"))
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(CUDA_SOURCES))) as ex:
        objs = list(ex.map(compile_one, CUDA_SOURCES))
    _run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(out), *map(str, objs)])
    stamp_file.write_text(stamp)
    if verbose:
        print(f"built {out}")
    return out


def build_host(force: bool = False, verbose: bool = False) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    out = host_lib_path()
    stamp_file = LIBDIR / ".host.stamp"
    srcs = [CSRC / s for s in HOST_SOURCES]
    stamp = _stamp(srcs + list((ROOT / "include").glob("*.h")))
    if out.exists() and not force and stamp_file.exists() and stamp_file.read_text() == stamp:
        return out
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    # -ffp-contract=off: the update arithmetic rounds exactly as written (no FMA contraction in the AVX2 clones)
    _run([cxx, "-O3", "-ffp-contract=off", "-std=c++17", "-fPIC", "-shared", "-pthread", "-o", str(out), *map(str, srcs), "-lrt"])
    stamp_file.write_text(stamp)
    if verbose:
        print(f"built {out}")
    return out


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_host(force, verbose)
    build_cuda(force, verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True)
