"""Single-node launcher: spawn N ranks on N GPUs (or N CPU processes).

Replaces the reference's ``mpirun --host ... python main.py ... distributed`` /
``horovodrun -np N ...`` command lines (fabfile.py:194-235) and its docker-compose two-node
"cluster" (docker-compose.yaml): on an NVSwitch box all ranks are local, so launching is
``fork + env``:

    python -m pytorch_distributed_rnn_b200.parallel.launch -n 8 -m pytorch_distributed_rnn_b200 --epochs 1 distributed
    python -m pytorch_distributed_rnn_b200.parallel.launch -n 2 examples/example_ddp.py

Each child gets RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR=127.0.0.1 / MASTER_PORT (the
variables torchrun sets, so scripts work under either).  Output is captured per rank and
returned/printed in rank order; a failing or timed-out rank kills the whole job (the
reference's only failure handling is a start timeout, fabfile.py:227).
"""
from __future__ import annotations

import os
import socket
import subprocess
import sys
import time
from typing import List, Optional, Sequence


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(cmd: Sequence[str], nproc: int, timeout: Optional[float] = None, env: Optional[dict] = None,
           capture: bool = True, master_port: Optional[int] = None,
           fail_fast: bool = True) -> List[subprocess.CompletedProcess]:
    """Run ``cmd`` as ``nproc`` ranks; returns one CompletedProcess per rank (rank order).  With ``fail_fast`` a
    rank that exits non-zero takes the others down (like mpirun); without it every rank runs to its own end."""
    port = master_port or free_port()
    procs = []
    for r in range(nproc):
        e = dict(os.environ)
        e.update(env or {})
        e.update(RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(nproc), LOCAL_WORLD_SIZE=str(nproc),
                 MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen(list(cmd), env=e, text=True,
                                      stdout=subprocess.PIPE if capture else None,
                                      stderr=subprocess.PIPE if capture else None))
    deadline = None if timeout is None else time.time() + timeout
    results: List[Optional[subprocess.CompletedProcess]] = [None] * nproc
    failed = False
    try:
        for r, p in enumerate(procs):
            left = None if deadline is None else max(0.1, deadline - time.time())
            try:
                out, err = p.communicate(timeout=left)
            except subprocess.TimeoutExpired:
                failed = True
                p.kill()
                out, err = p.communicate()
                err = (err or "") + f"\n[launch] rank {r} timed out"
            results[r] = subprocess.CompletedProcess(p.args, p.returncode, out, err)
            if p.returncode != 0 and fail_fast:
                failed = True
            if failed:
                for q in procs:
                    if q.poll() is None:
                        q.kill()
    finally:
        for q in procs:
            if q.poll() is None:
                q.kill()
    return results  # type: ignore[return-value]


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    n, timeout, module, rest = None, None, None, []
    i = 0
    while i < len(argv):  # own flags come first; everything from the script / module on is passed through
        a = argv[i]
        if a in ("-n", "--nproc"):
            n = None if argv[i + 1] == "auto" else int(argv[i + 1]); i += 2   # auto: one rank per visible GPU
        elif a == "--timeout":
            timeout = float(argv[i + 1]); i += 2
        elif a == "-m":
            module = argv[i + 1]; rest = argv[i + 2:]; break
        elif a in ("-h", "--help"):
            print(__doc__); return 0
        else:
            rest = argv[i:]; break
    if module is None and not rest:
        print(__doc__); return 2

    class _A:  # noqa: D401 - tiny namespace
        pass
    a = _A()
    a.module, a.rest, a.timeout = module, rest, timeout
    if n is None:
        try:
            import torch
            n = max(1, torch.cuda.device_count())
        except Exception:  # noqa: BLE001
            n = 1
    cmd = [sys.executable] + (["-m", a.module] if a.module else []) + a.rest
    res = launch(cmd, n, timeout=a.timeout)
    rc = 0
    for r, cp in enumerate(res):
        if cp.stdout:
            sys.stdout.write(cp.stdout)
        if cp.stderr:
            sys.stderr.write(cp.stderr)
        rc = rc or cp.returncode
    return rc


if __name__ == "__main__":
    raise SystemExit(main())
