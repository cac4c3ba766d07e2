"""Reference arm support (NOT product code): locating / installing the unmodified reference.

``ensure_reference()`` makes ``baseline/_ref/motion`` exist.  It is called by
``__graft_entry__.build()`` and by ``bench.py --impl reference`` so a fresh checkout gets a reference
arm without a manual step:

1. ``baseline/install_reference.sh`` — the offline ``pip install --target baseline/_ref`` of a /tmp copy
   of ``/root/reference`` (upstream has no setup.py / pyproject.toml, so a packaging-only shim is added
   to the copy; the installed sources are byte-identical to ``/root/reference/src``);
2. if pip is unusable: a byte-for-byte ``cp -r /root/reference/src/{motion,example} baseline/_ref/`` —
   the adapter only needs the files on ``sys.path``.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import time
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_ROOT = HERE / "_ref"
REF_SRC = Path(os.environ.get("B200RNN_REFERENCE_SRC", "/root/reference"))
_MARKER = REF_ROOT / "motion" / "trainer" / "base.py"


def reference_installed() -> bool:
    return _MARKER.exists()


def _install() -> str:
    if not (REF_SRC / "src" / "motion" / "trainer" / "base.py").exists():
        raise RuntimeError(f"reference sources not found at {REF_SRC} and baseline/_ref is empty")
    r = subprocess.run(["sh", str(HERE / "install_reference.sh")], capture_output=True, text=True)
    if r.returncode == 0 and reference_installed():
        return "pip --target (packaging shim on a /tmp copy)"
    REF_ROOT.mkdir(parents=True, exist_ok=True)
    for pkg in ("motion", "example"):
        dst = REF_ROOT / pkg
        if dst.exists():
            shutil.rmtree(dst)
        shutil.copytree(REF_SRC / "src" / pkg, dst)
    if not reference_installed():
        raise RuntimeError(f"could not populate {REF_ROOT}: pip said\n{r.stdout[-2000:]}\n{r.stderr[-2000:]}")
    return "byte-for-byte copy of src/{motion,example} (pip unusable)"


def ensure_reference(wait_s: float = 120.0) -> str:
    """Install the reference under baseline/_ref if it is missing.  Safe to call from every rank of a
    torchrun job: one process installs (lock directory), the others wait for the marker file."""
    if reference_installed():
        return "already installed"
    lock = HERE / "_ref.lock"
    try:
        if lock.exists() and time.time() - lock.stat().st_mtime > 600:
            lock.rmdir()      # left behind by an installer that died: nobody installs for ten minutes
    except OSError:
        pass
    try:
        lock.mkdir()
    except FileExistsError:
        deadline = time.time() + wait_s
        while not reference_installed():
            if time.time() > deadline:
                raise RuntimeError("timed out waiting for another process to install baseline/_ref")
            time.sleep(0.2)
        return "installed by another process"
    try:
        return _install()
    finally:
        try:
            lock.rmdir()
        except OSError:
            pass
