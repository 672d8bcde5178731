"""HW1 — distributed FP64 matrix multiply C = A*B (ref homeworks/hw1/src/template.c:30-238).

CLI as the reference: ``template <n>`` (default 64), n a power of two <= 4096 and divisible by the
number of processes; prints ``pid=0: ...`` lines, re-computes serially on the root, compares with
tolerance 1e-6 and prints ``Test: PASSED|FAILED``.

Two backends, same decomposition (1-D row blocks of A, B replicated):
* ``--backend cpu``  root sends row blocks of A (MPI_Send), broadcasts B, every rank runs the native
  triple loop, root receives the C blocks — over gloo (ref template.c:120-146).
* ``--backend gpu``  one process per GPU: A's block is uploaded locally, the FP64 tensor-core kernel
  (csrc/cuda/dgemm.cu) reads B straight from the root's device buffer and writes its C block straight
  into the root's result matrix over NVLink (peer memory) — no broadcast, no gather messages.
Default: gpu when there are at least as many GPUs as ranks, else cpu.
"""
from __future__ import annotations

import argparse
import ctypes as C
import sys
import time

import torch

from ..ops import cpu as cpu_ops
from ..parallel import dist as bdist

MAXDIM = 1 << 12
ROOT = 0


def _init_data(rows: int, cols: int) -> torch.Tensor:
    t = torch.empty((rows, cols), dtype=torch.float64)
    cpu_ops.lib().b200_hw1_init_data(C.c_void_p(t.data_ptr()), rows, cols)
    return t


def _matmul_cpu(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    rows, n = a.shape
    c = torch.empty((rows, n), dtype=torch.float64)
    cpu_ops.lib().b200_cpu_matmul_f64(C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()),
                                      C.c_void_p(c.data_ptr()), rows, n)
    return c


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="template")
    ap.add_argument("n", nargs="?", type=int, default=64)
    ap.add_argument("--backend", choices=["auto", "cpu", "gpu"], default="auto")
    ns = ap.parse_args(argv)
    rank, world, _ = bdist.env_rank_world()
    backend = ns.backend
    if backend == "auto":
        backend = "gpu" if torch.cuda.is_available() and torch.cuda.device_count() >= world else "cpu"
    ctx = bdist.init(device="cpu" if backend == "cpu" else None)
    n = ns.n
    if n <= 0 or (n & (n - 1)) != 0:
        if rank == ROOT:
            print(f"Error: Matrix dimension 'n' ({n}) must be a positive power of two.", file=sys.stderr)
        ctx.abort(1)
    if n > MAXDIM:
        if rank == ROOT:
            print(f"Warning: Matrix dimension 'n' ({n}) exceeds MAXDIM ({MAXDIM}). Clamping to MAXDIM.",
                  file=sys.stderr)
        n = MAXDIM
    if n % world != 0:
        if rank == ROOT:
            print(f"Error: Matrix dimension 'n' ({n}) must be divisible by the number of processes ({world}).",
                  file=sys.stderr)
        ctx.abort(1)
    my_work = n // world
    A = Bm = None
    if rank == ROOT:
        print(f"pid={rank}: num_procs={world} n={n} my_work={my_work} (rows per proc)", flush=True)
        A, Bm = _init_data(n, n), _init_data(n, n)

    ctx.barrier()
    t0 = time.perf_counter()
    if backend == "cpu":
        if rank == ROOT:
            reqs = [ctx.isend(A[p * my_work:(p + 1) * my_work], p, 0) for p in range(1, world)]
            local_a = A[:my_work]
            for r in reqs:
                r.wait()
        else:
            local_a = torch.empty((my_work, n), dtype=torch.float64)
            ctx.recv(local_a, ROOT, 0)
            Bm = torch.empty((n, n), dtype=torch.float64)
        ctx.bcast(Bm, ROOT)
        local_c = _matmul_cpu(local_a, Bm)
        if rank == ROOT:
            Cm = torch.empty((n, n), dtype=torch.float64)
            Cm[:my_work] = local_c
            for p in range(1, world):
                ctx.recv(Cm[p * my_work:(p + 1) * my_work], p, 0)
        else:
            ctx.send(local_c, ROOT, 0)
    else:
        from ..ops import cuda as cu
        from ..parallel.symm import IpcSymm, LocalSymm
        symm = IpcSymm(ctx) if world > 1 else LocalSymm(1, ctx.device).view(0)
        nb = n * n * 8
        bufB = symm.alloc("B", nb if rank == ROOT else 256)
        bufC = symm.alloc("C", nb if rank == ROOT else 256)
        symm.commit()
        dA = torch.empty((my_work, n), dtype=torch.float64, device=ctx.device)
        # A's row blocks still travel host->host (they start on the root's host, like the reference);
        # each rank uploads its own block over its own PCIe link.
        if rank == ROOT:
            reqs = [ctx.isend(A[p * my_work:(p + 1) * my_work], p, 0) for p in range(1, world)]
            bufB.tensor((n, n), torch.float64).copy_(Bm)
            dA.copy_(A[:my_work])
            for r in reqs:
                r.wait()
        else:
            la = torch.empty((my_work, n), dtype=torch.float64)
            ctx.recv(la, ROOT, 0)
            dA.copy_(la)
        torch.cuda.synchronize(ctx.device)
        ctx.barrier()                                   # B is resident on the root before anyone reads it
        cu.dgemm_rows(dA.data_ptr(), symm.ptr(ROOT, "B"),
                      symm.ptr(ROOT, "C") + rank * my_work * n * 8, my_work, n)
        torch.cuda.synchronize(ctx.device)
        ctx.barrier()                                   # every block of C has landed on the root
        if rank == ROOT:
            Cm = bufC.tensor((n, n), torch.float64).cpu()
    if rank == ROOT:
        elapsed = time.perf_counter() - t0
        print(f"pid={rank}: Parallel computation finished in {elapsed:f} seconds.")
        print(f"pid={rank}: Performing serial computation for verification...")
        s0 = time.perf_counter()
        D = _matmul_cpu(A, Bm)
        print(f"pid={rank}: Serial computation finished in {time.perf_counter() - s0:f} seconds.")
        flag = cpu_ops.lib().b200_hw1_check(C.c_void_p(Cm.data_ptr()), C.c_void_p(D.data_ptr()), n * n, 1e-6)
        print("--------------------------------------")
        if flag:
            bad = (Cm - D).abs().argmax().item()
            i, j = divmod(bad, n)
            print(f"ERROR: Mismatch found at C[{i}][{j}]={Cm[i, j]:f} != D[{i}][{j}]={D[i, j]:f} "
                  f"(Diff: {Cm[i, j] - D[i, j]:f})")
            print(f"pid={rank}: Test: FAILED")
        else:
            print(f"pid={rank}: Test: PASSED")
            print(f"pid={rank}: Total PARALLEL time: {elapsed:f} seconds.")
        print("--------------------------------------")
        print(f"backend={backend}", flush=True)
    ctx.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
