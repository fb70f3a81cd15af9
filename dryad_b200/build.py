"""In-tree native build: CUDA kernels (sm_100a) and the C++ host runtime.

`python -m dryad_b200.build` compiles every `csrc/kernels/*.cu` with
`nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo` into
`dryad_b200/_lib/libdryad_kernels.so` and every `csrc/host/*.cpp` into
`dryad_b200/_lib/libdryad_host.so`.  Objects are rebuilt only when a source or header is newer.
nvcc cross-compiles without a GPU, so this runs on the CPU-only build box; the `.so` files travel
with the repo snapshot to the GPU box.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
LIB = ROOT / "dryad_b200" / "_lib"
OBJ = ROOT / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v",
]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-pthread"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def _newer(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in deps)


def _run(cmd, log: Path | None = None) -> None:
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        log.write_text(p.stdout)
    if p.returncode != 0:
        sys.stderr.write(p.stdout)
        raise RuntimeError("build failed: " + " ".join(map(str, cmd)))


def build_kernels(force: bool = False, verbose: bool = False) -> Path:
    OBJ.mkdir(parents=True, exist_ok=True)
    LIB.mkdir(parents=True, exist_ok=True)
    srcs = sorted((CSRC / "kernels").glob("*.cu"))
    hdrs = list((CSRC / "kernels").glob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    nvcc = _nvcc()
    jobs = []
    for s in srcs:
        o = OBJ / (s.stem + ".o")
        if force or _newer(o, [s] + hdrs):
            jobs.append((s, o))

    def compile_one(job):
        s, o = job
        _run([nvcc, *NVCC_FLAGS, "-c", str(s), "-o", str(o)], OBJ / (s.stem + ".ptxas.log"))
        return s.name

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for name in ex.map(compile_one, jobs):
            if verbose:
                print("compiled", name)
    so = LIB / "libdryad_kernels.so"
    objs = [OBJ / (s.stem + ".o") for s in srcs]
    if force or jobs or _newer(so, objs):
        _run([nvcc, "-shared", "-o", str(so), *map(str, objs), "-lcuda"])
    return so


def build_host(force: bool = False, verbose: bool = False) -> Path | None:
    srcs = sorted((CSRC / "host").glob("*.cpp"))
    if not srcs:
        return None
    OBJ.mkdir(parents=True, exist_ok=True)
    LIB.mkdir(parents=True, exist_ok=True)
    hdrs = list((CSRC / "host").glob("*.h")) + list((ROOT / "include").glob("*.h"))
    cxx = os.environ.get("CXX", "g++")
    jobs = []
    for s in srcs:
        o = OBJ / ("host_" + s.stem + ".o")
        if force or _newer(o, [s] + hdrs):
            jobs.append((s, o))

    def compile_one(job):
        s, o = job
        _run([cxx, *CXX_FLAGS, "-I", str(ROOT / "include"), "-c", str(s), "-o", str(o)])
        return s.name

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for name in ex.map(compile_one, jobs):
            if verbose:
                print("compiled", name)
    so = LIB / "libdryad_host.so"
    objs = [OBJ / ("host_" + s.stem + ".o") for s in srcs]
    if force or jobs or _newer(so, objs):
        _run([cxx, "-shared", "-o", str(so), *map(str, objs), "-lz", "-pthread"])
    return so


def build_pyext(force: bool = False, verbose: bool = False) -> Path | None:
    """CPython extension modules (csrc/py/*.cpp -> dryad_b200/_lib/<name>.so): code that has to handle
    Python objects directly, such as the record codec.  Optional: without Python headers the pure-Python
    implementations are used."""
    import sysconfig
    srcs = sorted((CSRC / "py").glob("*.cpp"))
    inc = sysconfig.get_paths().get("include")
    if not srcs or not inc or not (Path(inc) / "Python.h").exists():
        return None
    LIB.mkdir(parents=True, exist_ok=True)
    cxx = os.environ.get("CXX", "g++")
    last = None
    for s in srcs:
        so = LIB / ("_dryad_" + s.stem.replace("py", "", 1) + ".so")
        if force or _newer(so, [s]):
            _run([cxx, *CXX_FLAGS, "-shared", "-I", inc, str(s), "-o", str(so)])
            if verbose:
                print("compiled", s.name)
        last = so
    return last


def build_all(force: bool = False, verbose: bool = False):
    k, h = build_kernels(force, verbose), build_host(force, verbose)
    build_pyext(force, verbose)
    return k, h


if __name__ == "__main__":
    k, h = build_all(force="--force" in sys.argv, verbose=True)
    print("built", k, h)
