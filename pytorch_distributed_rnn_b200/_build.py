"""In-tree build of the native library (``libb200rnn.so``) with nvcc for sm_100a.

No torch headers are involved: the library exposes a plain C ABI
(``include/b200rnn.h``) that the Python package drives through ``ctypes`` with raw
device pointers, so a full rebuild is seconds per file and cross-compiles on a box
without a GPU.  The reference ships no native code at all; this is the
"L0" it borrowed from torch/cuDNN/MPI, owned here.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO_ROOT = PKG_DIR.parent
CSRC = REPO_ROOT / "csrc"
INCLUDE = REPO_ROOT / "include"
BUILD_DIR = REPO_ROOT / "build" / "obj"
LIB_PATH = PKG_DIR / "_lib" / "libb200rnn.so"

NVCC_FLAGS = [
    "-std=c++17", "-O3",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
]
CXX_FLAGS = ["-std=c++17", "-O2", "-fPIC"]


def find_nvcc() -> str | None:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    return cand if cand and Path(cand).exists() else None


def _sources() -> list[Path]:
    return sorted(p for p in CSRC.rglob("*") if p.suffix in (".cu", ".cpp"))


def _headers_mtime() -> float:
    hs = list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.h")) + list(CSRC.rglob("*.inc")) + list(INCLUDE.rglob("*.h"))
    return max((h.stat().st_mtime for h in hs), default=0.0)


def is_stale() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    return any(s.stat().st_mtime > t for s in _sources()) or _headers_mtime() > t


def _compile_one(nvcc: str, src: Path, obj: Path, verbose: bool) -> str:
    cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-I", str(CSRC)]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    if src.suffix == ".cpp":
        cmd += ["-x", "cu"]  # host-only files still go through nvcc for the CUDA runtime headers
    cmd += ["-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return r.stderr


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every source under ``csrc/`` for sm_100a and link ``libb200rnn.so``."""
    if not force and not is_stale():
        return LIB_PATH
    nvcc = find_nvcc()
    if nvcc is None:
        raise RuntimeError("nvcc not found; cannot build libb200rnn.so")
    BUILD_DIR.mkdir(parents=True, exist_ok=True)
    LIB_PATH.parent.mkdir(parents=True, exist_ok=True)
    # one builder at a time: the ranks of a launch all import the package, and concurrent nvcc runs writing the
    # same object files corrupt each other.  The others wait for the lock and then find the library fresh.
    import fcntl
    with open(BUILD_DIR / ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        if not force and not is_stale():
            return LIB_PATH
        return _build_locked(nvcc, force, verbose)


def _build_locked(nvcc: str, force: bool, verbose: bool) -> Path:
    srcs = _sources()
    hdr_t = _headers_mtime()
    jobs = []
    objs = []
    for s in srcs:
        obj = BUILD_DIR / (str(s.relative_to(CSRC)).replace("/", "__") + ".o")
        objs.append(obj)
        if force or not obj.exists() or obj.stat().st_mtime < max(s.stat().st_mtime, hdr_t):
            jobs.append((s, obj))
    logs = []
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for log in ex.map(lambda so: _compile_one(nvcc, so[0], so[1], verbose), jobs):
                logs.append(log)
    tmp = LIB_PATH.with_suffix(".so.tmp%d" % os.getpid())
    link = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a",
            "-o", str(tmp), *map(str, objs), "-lpthread", "-ldl"]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, LIB_PATH)
    if verbose:
        sys.stderr.write("\n".join(logs))
    return LIB_PATH


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(p)
