"""In-tree native builds (no JIT cache): ``lib/libb200cnn_cpu.so`` with g++ and
``lib/libb200cnn_cuda.so`` with nvcc for sm_100a only.

The .so files are git-ignored but travel with the gpurun snapshot; both are rebuilt when a
source or header is newer than the library.  This replaces the reference's per-version
Makefiles' compile step (ref final_project/v3_cuda_only/Makefile:6-16,
v4_mpi_cuda/Makefile:11-42) for the framework libraries; the per-version ``template``
launchers under ``final_project/`` still expose ``make GPU_ARCH_FLAGS=...``.
"""
from __future__ import annotations

import contextlib
import fcntl
import glob
import os
import shutil
import subprocess
import sys

from .. import REPO_ROOT

LIB_DIR = os.path.join(REPO_ROOT, "lib")
INC_DIR = os.path.join(REPO_ROOT, "include")
CPU_SRC = sorted(glob.glob(os.path.join(REPO_ROOT, "csrc", "cpu", "*.cpp")))
CPU_LIB = os.path.join(LIB_DIR, "libb200cnn_cpu.so")
CUDA_LIB = os.path.join(LIB_DIR, "libb200cnn_cuda.so")

SM100A = ["-gencode", "arch=compute_100a,code=sm_100a"]


def cuda_sources():
    return sorted(glob.glob(os.path.join(REPO_ROOT, "csrc", "cuda", "*.cu")))


def _headers():
    return glob.glob(os.path.join(INC_DIR, "*.h")) + \
        glob.glob(os.path.join(REPO_ROOT, "csrc", "cuda", "*.cuh"))


def _stale(lib: str, srcs) -> bool:
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    return any(os.path.getmtime(s) > t for s in list(srcs) + _headers())


def _run(cmd, what):
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if p.returncode != 0:
        sys.stderr.write(p.stdout)
        raise RuntimeError(f"building {what} failed: {' '.join(cmd)}")
    return p.stdout


@contextlib.contextmanager
def _build_lock():
    """One builder at a time per checkout: every rank of a torchrun / bin/mpirun job imports this module, and
    without the lock they would all run nvcc into the same lib/obj/*.o and link half-written objects.  The
    staleness test is repeated under the lock, so the ranks that waited find a fresh library and build nothing."""
    os.makedirs(LIB_DIR, exist_ok=True)
    with open(os.path.join(LIB_DIR, ".build.lock"), "w") as f:
        fcntl.flock(f, fcntl.LOCK_EX)
        try:
            yield
        finally:
            fcntl.flock(f, fcntl.LOCK_UN)


CPU_WARN = ["-Wall", "-Wextra", "-Werror"]          # the CPU library builds warning-free; keep it that way


def build_cpu(force: bool = False) -> str:
    srcs = sorted(glob.glob(os.path.join(REPO_ROOT, "csrc", "cpu", "*.cpp")))
    if not (force or _stale(CPU_LIB, srcs)):
        return CPU_LIB
    with _build_lock():
        if force or _stale(CPU_LIB, srcs):
            tmp = CPU_LIB + f".tmp{os.getpid()}"
            _run(["g++", "-std=c++17", "-O3", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared"] + CPU_WARN +
                 ["-I", INC_DIR, "-o", tmp] + srcs, "CPU library")
            os.replace(tmp, CPU_LIB)
    return CPU_LIB


def build_cpu_sanitized(out_dir: str) -> str:
    """``libb200cnn_cpu.so`` with AddressSanitizer + UndefinedBehaviourSanitizer (scripts/sanitize_host.sh loads it
    into a test run with LD_PRELOAD=libasan): the host-side counterpart of compute-sanitizer, which is closed on
    the GPU pool.  The reference ships a .clang-tidy it never runs (ref final_project/v4_mpi_cuda/.clang-tidy:4)."""
    srcs = sorted(glob.glob(os.path.join(REPO_ROOT, "csrc", "cpu", "*.cpp")))
    os.makedirs(out_dir, exist_ok=True)
    lib = os.path.join(out_dir, "libb200cnn_cpu.so")
    _run(["g++", "-std=c++17", "-O1", "-g", "-fno-omit-frame-pointer", "-ffp-contract=off", "-fopenmp", "-fPIC",
          "-shared", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"] + CPU_WARN +
         ["-I", INC_DIR, "-o", lib] + srcs, "sanitized CPU library")
    return lib


def nvcc_path():
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    return None


def build_cuda(force: bool = False, verbose: bool = False) -> str:
    """nvcc cross-compiles sm_100a without a GPU.  One translation unit per .cu, linked into
    one shared object; -lineinfo keeps ncu's source page usable."""
    srcs = cuda_sources()
    if not (force or _stale(CUDA_LIB, srcs)):
        return CUDA_LIB
    with _build_lock():
        return _build_cuda_locked(srcs, force, verbose)


def _build_cuda_locked(srcs, force: bool, verbose: bool) -> str:
    nvcc = nvcc_path()
    if not (force or _stale(CUDA_LIB, srcs)):         # another rank built it while we waited for the lock
        return CUDA_LIB
    if nvcc is None:
        if os.path.exists(CUDA_LIB):
            return CUDA_LIB       # stale but usable (GPU box without sources changed)
        raise RuntimeError("nvcc not found and libb200cnn_cuda.so is not built")
    os.makedirs(LIB_DIR, exist_ok=True)
    objdir = os.path.join(LIB_DIR, "obj")
    os.makedirs(objdir, exist_ok=True)
    objs, procs = [], []
    for s in srcs:
        o = os.path.join(objdir, os.path.basename(s) + ".o")
        objs.append(o)
        if not force and os.path.exists(o) and not _stale(o, [s]):
            continue
        cmd = [nvcc, "-std=c++17", "-O3", "-lineinfo"] + SM100A + [
            "-Xcompiler", "-fPIC", "-Xptxas", "-v" if verbose else "-O3",
            "-I", INC_DIR, "-c", s, "-o", o]
        procs.append((s, cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE,
                                               stderr=subprocess.STDOUT, text=True)))
    for s, cmd, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError(f"nvcc failed on {s}: {' '.join(cmd)}")
        if verbose:
            sys.stderr.write(out)
    tmp = CUDA_LIB + f".tmp{os.getpid()}"
    _run([nvcc, "-shared"] + SM100A + ["-o", tmp] + objs, "CUDA library")
    os.replace(tmp, CUDA_LIB)
    return CUDA_LIB


if __name__ == "__main__":
    print(build_cpu(force="--force" in sys.argv))
    print(build_cuda(force="--force" in sys.argv, verbose="-v" in sys.argv))
