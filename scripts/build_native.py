#!/usr/bin/env python
"""Builds the native libraries in-tree.

  libdfm_cuda.so  csrc/cuda/*.cu   nvcc, sm_100a only (cross-compiles without a GPU)
  libdfm_io.so    csrc/io/*.cc     g++ -O3, host-only data layer
  libdfm_cpu.so   csrc/cpu/*.cc    g++ -O3, host compute ops of the CPU trainers (AVX2 clones picked at load time)

Objects are rebuilt only when a source or header is newer than the object.
Outputs land in <package>/_native/ so they travel with the source snapshot.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "deepfm-tensorflow-distributed-training-on-amazon-sagemaker_b200"
OUT = PKG / "_native"
OBJ = ROOT / "build" / "obj"

NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
CXX = os.environ.get("CXX") or shutil.which("g++") or "g++"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    "-I", str(ROOT / "include"),
]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-pthread", "-msse4.2", "-Wall", "-I", str(ROOT / "include")]


def _newer(src_list, target: Path) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(s).stat().st_mtime > t for s in src_list)


def _run(cmd, log: Path | None = None):
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        log.write_text(p.stdout)
    if p.returncode != 0:
        sys.stderr.write(" ".join(map(str, cmd)) + "\n" + p.stdout + "\n")
        raise RuntimeError(f"build step failed: {cmd[0]} {cmd[-1]}")
    return p.stdout


def build_cuda(verbose: bool = False) -> Path:
    srcs = sorted((ROOT / "csrc" / "cuda").glob("*.cu"))
    hdrs = list((ROOT / "csrc" / "cuda").glob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    OBJ.mkdir(parents=True, exist_ok=True)
    OUT.mkdir(parents=True, exist_ok=True)
    lib = OUT / "libdfm_cuda.so"
    jobs = []
    for s in srcs:
        o = OBJ / (s.stem + ".o")
        if _newer([s] + hdrs, o):
            jobs.append((s, o))
    def one(job):
        s, o = job
        out = _run([NVCC, *NVCC_FLAGS, "-c", str(s), "-o", str(o)], OBJ / (s.stem + ".ptxas.log"))
        if verbose:
            print(out)
    with cf.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(one, jobs))
    objs = [OBJ / (s.stem + ".o") for s in srcs]
    if jobs or _newer(objs, lib):
        _run([NVCC, "-shared", "-o", str(lib), *map(str, objs), "-cudart", "shared",
              "-gencode", "arch=compute_100a,code=sm_100a"])
    return lib


def build_io(verbose: bool = False) -> Path:
    srcs = sorted((ROOT / "csrc" / "io").glob("*.cc"))
    hdrs = list((ROOT / "csrc" / "io").glob("*.h")) + list((ROOT / "include").glob("*.h"))
    OUT.mkdir(parents=True, exist_ok=True)
    lib = OUT / "libdfm_io.so"
    if srcs and _newer(srcs + hdrs, lib):
        _run([CXX, *CXX_FLAGS, "-shared", "-o", str(lib), *map(str, srcs)])
    return lib


def build_cpu(verbose: bool = False) -> Path:
    srcs = sorted((ROOT / "csrc" / "cpu").glob("*.cc"))
    OUT.mkdir(parents=True, exist_ok=True)
    lib = OUT / "libdfm_cpu.so"
    hdrs = list((ROOT / "csrc" / "cpu").glob("*.h"))
    if srcs and _newer(srcs + hdrs, lib):
        _run([CXX, "-O3", "-std=c++17", "-fPIC", "-pthread", "-Wall", "-shared", "-o", str(lib), *map(str, srcs)])
    return lib


def main():
    verbose = "-v" in sys.argv
    what = [a for a in sys.argv[1:] if not a.startswith("-")] or ["cuda", "io", "cpu"]
    if "cuda" in what:
        print("built", build_cuda(verbose))
    if "io" in what:
        print("built", build_io(verbose))
    if "cpu" in what:
        print("built", build_cpu(verbose))


if __name__ == "__main__":
    main()
