"""Python face of the native bootstrap (csrc/runtime/bootstrap.cpp): rendezvous + topology.

``NativeBootstrap`` is a star of TCP connections to rank 0 carrying small control messages
(CUDA IPC handles, barriers).  It plays the role ``mpirun`` + ``MPI_Init`` play for the
reference (src/motion/trainer/ddp.py:18) without needing MPI, ssh keys or a hostfile: on an
NVSwitch node all ranks are local processes, one per GPU.  It is pure host code and is tested
with two CPU processes.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

from .. import _native


class NativeBootstrap:
    def __init__(self, rank: int, world: int, addr: str = "127.0.0.1", port: Optional[int] = None,
                 listen_fd: int = -1, timeout_ms: int = 60000):
        self.lib = _native.lib()
        self.rank, self.world = rank, world
        self.handle = C.c_void_p()
        if port is None:
            port = int(os.environ.get("B200RNN_BOOT_PORT", "0"))
        if rank == 0 and listen_fd < 0 and world > 1:
            port, listen_fd = self.listen(port)
        self.port = port
        _native.check(self.lib.b200rnn_boot_init(rank, world, addr.encode(), port, listen_fd, timeout_ms,
                                                 C.byref(self.handle)), "boot_init")

    @staticmethod
    def listen(port: int = 0):
        lib = _native.lib()
        p, fd = C.c_int32(port), C.c_int32(-1)
        _native.check(lib.b200rnn_boot_listen(C.byref(p), C.byref(fd)), "boot_listen")
        return p.value, fd.value

    def allgather(self, payload: bytes) -> List[bytes]:
        n = len(payload)
        out = C.create_string_buffer(n * self.world)
        _native.check(self.lib.b200rnn_boot_allgather(self.handle, payload, n, out), "boot_allgather")
        return [out.raw[i * n:(i + 1) * n] for i in range(self.world)]

    def barrier(self) -> None:
        _native.check(self.lib.b200rnn_boot_barrier(self.handle), "boot_barrier")

    def close(self) -> None:
        if self.handle:
            self.lib.b200rnn_boot_destroy(self.handle)
            self.handle = C.c_void_p()


def probe_topology() -> dict:
    """NVLink / P2P capabilities of the local GPUs (cudaDeviceCanAccessPeer, P2P attributes, NVLS)."""
    lib = _native.lib()
    t = _native.Topology()
    _native.check(lib.b200rnn_topology_probe(C.byref(t)), "topology_probe")
    n = t.n_devices
    return {
        "n_devices": n,
        "names": [t.name[i].value.decode() for i in range(n)],
        "sm_count": [t.sm_count[i] for i in range(n)],
        "multicast_supported": [bool(t.multicast_supported[i]) for i in range(n)],
        "can_access": [[int(t.can_access[i][j]) for j in range(n)] for i in range(n)],
        "perf_rank": [[int(t.perf_rank[i][j]) for j in range(n)] for i in range(n)],
        "native_atomics": [[int(t.native_atomics[i][j]) for j in range(n)] for i in range(n)],
        "all_to_all": all(t.can_access[i][j] for i in range(n) for j in range(n)),
    }
