"""Builds libpdcc.so in-tree with nvcc for sm_100a (no JIT cache, no torch ABI).

One nvcc process per translation unit, run in parallel, then one link step. The
result lands next to this file so that it travels with a `gpurun` snapshot.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
CSRC = ROOT / "csrc"
OBJ = HERE / "obj"
LIB = HERE / "libpdcc.so"

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall",
          "-I", str(ROOT / "include"), "-I", str(CSRC)]


def _nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(nvcc).exists():
        raise RuntimeError("nvcc not found; cannot build libpdcc.so")
    return nvcc


def sources() -> list[Path]:
    return sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cpp")))


def _digest() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.iterdir()) + [ROOT / "include" / "pdcc.h", Path(__file__)]):
        if p.is_file():
            h.update(p.name.encode())
            h.update(p.read_bytes())
    return h.hexdigest()


def _compile(src: Path, verbose: bool) -> Path:
    obj = OBJ / (src.name + ".o")
    cmd = [_nvcc(), *ARCH, *COMMON, "-c", str(src), "-o", str(obj)]
    if src.suffix == ".cu":
        cmd[1:1] = ["-Xptxas", "-v"] if verbose else []
    else:
        cmd[1:1] = ["-x", "cu"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src.name}:\n{r.stdout}\n{r.stderr}")
    if verbose:
        (OBJ / (src.name + ".ptxas.txt")).write_text(r.stderr)
    return obj


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every CUDA source for sm_100a and link libpdcc.so. Returns its path."""
    stamp = HERE / "libpdcc.digest"
    dig = _digest()
    if not force and LIB.exists() and stamp.exists() and stamp.read_text() == dig:
        return LIB
    OBJ.mkdir(exist_ok=True)
    srcs = sources()
    with cf.ThreadPoolExecutor(max_workers=min(len(srcs), os.cpu_count() or 4)) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), srcs))
    cmd = [_nvcc(), *ARCH, "-shared", "-o", str(LIB), *map(str, objs), "-lpthread", "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(dig)
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv or "-v" in sys.argv)
    print(path)
