"""Build libla_b200.so (sm_100a only) in-tree with nvcc.

    python -m linalg_b200.build [--force] [--verbose]

Every .cu under linalg_b200/csrc is compiled to an object with
`-gencode arch=compute_100a,code=sm_100a -lineinfo` and linked into
linalg_b200/lib/libla_b200.so (static cudart, so the library is self-contained and
travels to the GPU box with the repo snapshot).  Objects are rebuilt only when the
source or a header is newer.
"""
from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
CSRC = ROOT / "csrc"
OBJ = ROOT / "build"
LIBDIR = ROOT / "lib"
LIB = LIBDIR / "libla_b200.so"

NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CFLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
          "-Xptxas", "-v" if os.environ.get("LA_PTXAS_V") else "-O3"]


def _sources():
    return sorted(CSRC.glob("*.cu"))


def _headers_mtime():
    return max((p.stat().st_mtime for p in list(CSRC.glob("*.h")) + list(CSRC.glob("*.cuh"))),
               default=0.0)


def _defines():
    return os.environ.get("LA_EXTRA_DEFINES", "").split()


def _compile(src: Path, force: bool, verbose: bool, hdr_mtime: float, defines) -> Path:
    obj = OBJ / (src.stem + ".o")
    stamp = OBJ / (src.stem + ".flags")
    flags = " ".join(ARCH + CFLAGS + defines)
    if (not force and obj.exists() and obj.stat().st_mtime > max(src.stat().st_mtime, hdr_mtime)
            and stamp.exists() and stamp.read_text() == flags):
        return obj
    cmd = [NVCC, *ARCH, *CFLAGS, *defines, "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError(f"nvcc failed on {src.name}")
    if verbose or os.environ.get("LA_PTXAS_V"):
        sys.stderr.write(r.stderr)
    stamp.write_text(flags)
    return obj


def build(force: bool = False, verbose: bool = False) -> Path:
    OBJ.mkdir(exist_ok=True)
    LIBDIR.mkdir(exist_ok=True)
    srcs = _sources()
    hdr = _headers_mtime()
    defines = _defines()
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(lambda s: _compile(s, force, verbose, hdr, defines), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if force or not LIB.exists() or LIB.stat().st_mtime < newest:
        cmd = [NVCC, *ARCH, "-shared", "-o", str(LIB), *map(str, objs), "-Xlinker", "-Bsymbolic",
               "-Xlinker", "--no-undefined", "-lcuda" if False else "-ldl"]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return LIB


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(p)
