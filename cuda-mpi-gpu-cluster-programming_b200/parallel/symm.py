"""Symmetric device memory: named buffers every rank allocates and every rank can address.

Replaces the host ``std::vector`` message buffers of the reference.  Two
implementations behind one interface (``alloc`` / ``commit`` / ``ptr``):

* :class:`IpcSymm` — one process per GPU: ``cudaMalloc`` + CUDA IPC handles, exchanged once via
  ``torch.distributed`` (object all-gather on the bootstrap group); kernels then dereference the
  peers' buffers directly over NVLink.
* :class:`LocalSymm` — all ranks inside one process on one device (single-GPU emulation of the
  multi-rank protocol for tests; B200_PROFILING.md forbids spinning co-scheduled ranks on one
  GPU, so the emulation launches layer by layer and no wait ever blocks).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Tuple

import torch

from ..ops import cuda as cu


class DevBuf:
    """Raw device allocation viewed zero-copy as torch tensors via __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int, device: torch.device, owner=None):
        self.ptr, self.nbytes, self.device, self._owner = ptr, nbytes, device, owner

    def tensor(self, shape, dtype=torch.float32, offset_bytes: int = 0) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= s
        item = torch.empty((), dtype=dtype).element_size()
        assert offset_bytes + n * item <= self.nbytes, "view exceeds allocation"
        typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1",
                   torch.float64: "<f8"}[dtype]

        class _V:
            pass
        v = _V()
        v.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr,
                                      "data": (self.ptr + offset_bytes, False), "version": 3,
                                      "strides": None}
        v._keep = self
        t = torch.as_tensor(v, device=self.device)
        t._b200_keep = self
        return t


class _OwnedAlloc:
    def __init__(self, ptr):
        self.ptr = ptr

    def __del__(self):
        try:
            if self.ptr:
                cu.lib().b200_symm_free(C.c_void_p(self.ptr))
        except Exception:
            pass


class IpcSymm:
    """One process per GPU.  ``dist`` is a :class:`b200cnn.parallel.dist.DistContext`."""

    def __init__(self, dist):
        self.dist = dist
        self.rank, self.world = dist.rank, dist.world
        self.device = dist.device
        self._mine: Dict[str, DevBuf] = {}
        self._peer: Dict[Tuple[int, str], int] = {}
        self._opened: List[int] = []

    def alloc(self, name: str, nbytes: int) -> DevBuf:
        nbytes = max(int(nbytes), 256)
        p = C.c_void_p()
        with torch.cuda.device(self.device):
            cu.cu_check(cu.lib().b200_symm_alloc(nbytes, C.byref(p)), f"symm_alloc({name})")
        buf = DevBuf(p.value, nbytes, self.device, _OwnedAlloc(p.value))
        self._mine[name] = buf
        return buf

    def commit(self) -> None:
        """Exchange IPC handles for every buffer allocated so far and map the peers'."""
        mine = {}
        with torch.cuda.device(self.device):
            for name, b in self._mine.items():
                if (self.rank, name) in self._peer:
                    continue
                h = C.create_string_buffer(64)
                cu.cu_check(cu.lib().b200_ipc_export(C.c_void_p(b.ptr), h), f"ipc_export({name})")
                mine[name] = (h.raw, b.nbytes)
                self._peer[(self.rank, name)] = b.ptr
            allh = self.dist.all_gather_object(mine)
            for r, table in enumerate(allh):
                if r == self.rank:
                    continue
                for name, (raw, _n) in table.items():
                    p = C.c_void_p()
                    cu.cu_check(cu.lib().b200_ipc_open(raw, C.byref(p)), f"ipc_open({name}@{r})")
                    self._peer[(r, name)] = p.value
                    self._opened.append(p.value)
        self.dist.barrier()

    def ptr(self, rank: int, name: str) -> int:
        return self._peer[(rank, name)]

    def has(self, rank: int, name: str) -> bool:
        return (rank, name) in self._peer

    def close(self):
        for p in self._opened:
            cu.lib().b200_ipc_close(C.c_void_p(p))
        self._opened.clear()


class LocalSymm:
    """All ranks in this process, all buffers on one device."""

    def __init__(self, world: int, device):
        self.world = world
        self.device = torch.device(device)
        self._bufs: Dict[Tuple[int, str], DevBuf] = {}

    def view(self, rank: int) -> "_LocalView":
        return _LocalView(self, rank)


class _LocalView:
    def __init__(self, parent: LocalSymm, rank: int):
        self.parent, self.rank, self.world, self.device = parent, rank, parent.world, parent.device

    def alloc(self, name: str, nbytes: int) -> DevBuf:
        nbytes = max(int(nbytes), 256)
        t = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
        buf = DevBuf(t.data_ptr(), nbytes, self.device, t)
        self.parent._bufs[(self.rank, name)] = buf
        return buf

    def commit(self) -> None:
        pass

    def ptr(self, rank: int, name: str) -> int:
        return self.parent._bufs[(rank, name)].ptr

    def has(self, rank: int, name: str) -> bool:
        return (rank, name) in self.parent._bufs

    def close(self):
        pass
