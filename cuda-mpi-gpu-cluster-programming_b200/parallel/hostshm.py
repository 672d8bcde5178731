"""One pinned host tensor that every rank of a node writes into.

The reference gathers the result with a D2H copy per rank, a host-side trim and ``MPI_Gatherv`` into
rank 0's ``std::vector`` (ref final_project/v4_mpi_cuda/src/main_mpi_cuda.cpp:100-130).  Here rank 0 creates a
POSIX shared-memory file, every rank maps it and pins it (``cudaHostRegister``), and each rank's copy engine
writes the rows it owns straight to their final place over that rank's own PCIe link — no hop through
rank 0's HBM and no single link that carries everybody's bytes.  A few ints at the end of the segment are
per-rank "step s has landed" flags, written by a one-thread kernel in stream order after the copy and polled
by whoever consumes the result.
"""
from __future__ import annotations

import mmap
import os
import time
from typing import Optional

import numpy as np
import torch

from ..ops import cuda as cu

_FLAG_INTS = 64          # flag block: one int per (slot, rank), padded


class SharedHostTensor:
    """``slots`` tensors of ``shape`` (float32) + a flag block, shared by all ranks of ``dist``
    (``dist=None``: a private, process-local instance with the same interface)."""

    def __init__(self, shape, slots: int, dist=None, device=None):
        self.shape = tuple(int(s) for s in shape)
        self.slots = slots
        self.world = dist.world if dist is not None else 1
        self.rank = dist.rank if dist is not None else 0
        n = 1
        for s in self.shape:
            n *= s
        self.slot_bytes = (n * 4 + 4095) // 4096 * 4096
        nflag = max(_FLAG_INTS, slots * self.world)
        self.flag_off = slots * self.slot_bytes
        self.nbytes = self.flag_off + (nflag * 4 + 4095) // 4096 * 4096
        name = None
        if self.rank == 0:
            name = f"/dev/shm/b200cnn_{os.getpid()}_{int(time.time() * 1e6) & 0xFFFFFF:x}"
            with open(name, "wb") as f:
                f.truncate(self.nbytes)
        if dist is not None:
            name = dist.broadcast_object(name, 0)
        self.path = name
        fd = os.open(name, os.O_RDWR)
        try:
            self._mm = mmap.mmap(fd, self.nbytes, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
        finally:
            os.close(fd)
        self._np = np.frombuffer(self._mm, dtype=np.uint8)
        self.host_ptr = self._np.ctypes.data
        with torch.cuda.device(device if device is not None else (dist.device if dist else "cuda:0")):
            self.dev_ptr = cu.host_register(self.host_ptr, self.nbytes)
        if dist is not None:
            dist.barrier()                                            # everyone has it mapped: the name can go
        if self.rank == 0:
            try:
                os.unlink(name)
            except OSError:
                pass
        self._flags = np.frombuffer(self._mm, dtype=np.int32, count=nflag, offset=self.flag_off)
        self.views = []
        for i in range(slots):
            a = np.frombuffer(self._mm, dtype=np.float32, count=n, offset=i * self.slot_bytes)
            self.views.append(torch.from_numpy(a).view(self.shape))

    # -- addresses -------------------------------------------------------------------------------
    def slot_host_ptr(self, slot: int) -> int:
        return self.host_ptr + slot * self.slot_bytes

    def flag_dev_ptr(self, slot: int, rank: Optional[int] = None) -> int:
        r = self.rank if rank is None else rank
        return self.dev_ptr + self.flag_off + 4 * (slot * self.world + r)

    def flag(self, slot: int, rank: int) -> int:
        return int(self._flags[slot * self.world + rank])

    def wait_all(self, slot: int, value: int, ranks, timeout_s: float = 60.0) -> None:
        """Spin (host) until every rank in ``ranks`` has published ``>= value`` for ``slot``."""
        t0 = time.perf_counter()
        idx = [slot * self.world + r for r in ranks]
        spins = 0
        while True:
            if all(int(self._flags[i]) >= value for i in idx):
                return
            spins += 1
            if (spins & 0x3FF) == 0 and time.perf_counter() - t0 > timeout_s:
                raise TimeoutError(f"host gather: slot {slot} never reached step {value}: "
                                   f"{[int(self._flags[i]) for i in idx]}")

    def close(self) -> None:
        try:
            cu.host_unregister(self.host_ptr)
        except Exception:
            pass
