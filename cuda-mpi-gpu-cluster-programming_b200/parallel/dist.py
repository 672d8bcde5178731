"""Rank bootstrap over ``torch.distributed`` — the framework's MPI_Init/Comm_rank/Comm_size.

The reference's entire backend API is 18 MPI calls on MPI_COMM_WORLD.  Here
torch.distributed does the *plumbing only*: rendezvous, the small object exchanges (IPC handles),
host-side barriers, and the CPU (gloo) message path of the V2 runners.  Data on the GPU path
never goes through it — kernels use peer memory (parallel/symm.py).

Environment: torchrun's RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT, or Open MPI's
OMPI_COMM_WORLD_* (so ``bin/mpirun -np N`` works like the reference's launch line,
ref scripts/common_test_utils.sh:273-276).
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass
from typing import Any, List, Optional

import torch


def env_rank_world():
    e = os.environ
    rank = int(e.get("RANK", e.get("OMPI_COMM_WORLD_RANK", e.get("PMI_RANK", "0"))))
    world = int(e.get("WORLD_SIZE", e.get("OMPI_COMM_WORLD_SIZE", e.get("PMI_SIZE", "1"))))
    local = int(e.get("LOCAL_RANK", e.get("OMPI_COMM_WORLD_LOCAL_RANK", str(rank))))
    return rank, world, local


@dataclass
class DistContext:
    rank: int
    world: int
    local_rank: int
    device: torch.device
    backend: str
    initialized_here: bool = False

    @property
    def is_root(self) -> bool:
        return self.rank == 0

    # -- collectives used for plumbing -------------------------------------------------------
    def barrier(self) -> None:
        if self.world > 1:
            import torch.distributed as td
            td.barrier()

    def all_gather_object(self, obj: Any) -> List[Any]:
        if self.world == 1:
            return [obj]
        import torch.distributed as td
        out = [None] * self.world
        td.all_gather_object(out, obj)
        return out

    def broadcast_object(self, obj: Any, src: int = 0) -> Any:
        if self.world == 1:
            return obj
        import torch.distributed as td
        box = [obj]
        td.broadcast_object_list(box, src=src)
        return box[0]

    def max_float(self, v: float) -> float:
        """max over ranks (every multi-GPU time is reported as the slowest rank's)."""
        if self.world == 1:
            return float(v)
        return max(self.all_gather_object(float(v)))

    # -- CPU message path (V2 runners): MPI_Send/Recv/Bcast equivalents over gloo -------------
    def send(self, t: torch.Tensor, dst: int, tag: int = 0) -> None:
        import torch.distributed as td
        td.send(t.contiguous(), dst=dst, tag=tag)

    def recv(self, t: torch.Tensor, src: int, tag: int = 0) -> None:
        import torch.distributed as td
        td.recv(t, src=src, tag=tag)

    def isend(self, t: torch.Tensor, dst: int, tag: int = 0):
        import torch.distributed as td
        return td.isend(t.contiguous(), dst=dst, tag=tag)

    def irecv(self, t: torch.Tensor, src: int, tag: int = 0):
        import torch.distributed as td
        return td.irecv(t, src=src, tag=tag)

    def bcast(self, t: torch.Tensor, src: int = 0) -> None:
        if self.world > 1:
            import torch.distributed as td
            td.broadcast(t, src=src)

    def abort(self, code: int = 1) -> None:
        """MPI_Abort equivalent: fail-stop for every rank."""
        import sys
        sys.stderr.write(f"[Rank {self.rank}] aborting with code {code}\n")
        sys.stderr.flush()
        os._exit(code)

    def finalize(self) -> None:
        if self.initialized_here and self.world > 1:
            import torch.distributed as td
            if td.is_initialized():
                try:
                    td.barrier()
                except Exception:
                    pass
                td.destroy_process_group()


def init(device: Optional[str] = None, timeout_s: int = 600) -> DistContext:
    """Bootstraps ranks.  ``device='cpu'`` forces the gloo-only CPU plumbing."""
    rank, world, local = env_rank_world()
    use_cuda = (device != "cpu") and torch.cuda.is_available()
    if use_cuda:
        ndev = torch.cuda.device_count()
        # One GPU per rank: the device-side waits of the multi-GPU forward spin on peers' flags and would
        # deadlock (until the watchdog) if two ranks shared a device.  More ranks than GPUs is the job of the
        # single-device emulation (runners/v4.py::_oversubscribed), never of a silent ``local % ndev``.
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("OMPI_COMM_WORLD_LOCAL_SIZE", world)))
        if world > 1 and local_world > ndev and os.environ.get("B200_ALLOW_SHARED_GPU") != "1":
            raise RuntimeError(f"[Rank {rank}] {local_world} ranks on this node but only {ndev} visible GPU(s): ranks "
                               f"that wait on each other's device flags must not share a GPU (run the emulated path, "
                               f"e.g. runners.v4 does it automatically, or set CUDA_VISIBLE_DEVICES / -np accordingly)")
        dev = torch.device("cuda", local % max(ndev, 1))
        torch.cuda.set_device(dev)
    else:
        dev = torch.device("cpu")
    here = False
    backend = "none"
    if world > 1:
        import torch.distributed as td
        if not td.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            backend = "cpu:gloo,cuda:nccl" if use_cuda else "gloo"
            kw = dict(backend=backend, rank=rank, world_size=world,
                      timeout=datetime.timedelta(seconds=timeout_s))
            if use_cuda:
                kw["device_id"] = dev
            td.init_process_group(**kw)
            here = True
        else:
            backend = str(td.get_backend())
    return DistContext(rank, world, local, dev, backend, here)
