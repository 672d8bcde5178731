#!/usr/bin/env python3
"""Build the UNMODIFIED reference into baseline/_ref/ (git-ignored) for same-box baselines.

The prescribed `pip install --target baseline/_ref /root/reference` cannot work: the reference is
a C++/CUDA course repository with no setup.py / pyproject.toml (pip: "Directory is not
installable").  So its sources are copied verbatim and built with their own Makefiles / flags:

  v1   g++ -std=c++11 -O3                                  (ref final_project/v1_serial/Makefile)
  v2.1, v2.2  mpicxx -O3  -> baseline/bin/mpicxx (g++ + baseline/mpishim: the image has no MPI)
  v3   make GPU_ARCH_FLAGS="-gencode arch=compute_100,code=sm_100"     (ref v3 Makefile:6)
  v4   same flags, nvcc -ccbin=mpicxx                                   (ref v4 Makefile:26)
  hw1  mpicc

Nothing under baseline/_ref/src is edited.  `--check` only reports what is built.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("B200_REFERENCE", "/root/reference")
DST = os.path.join(HERE, "_ref")
SM100 = "-gencode arch=compute_100,code=sm_100"


def sh(cmd, cwd, log):
    with open(log, "a") as f:
        f.write(f"$ {cmd}\n")
        f.flush()
        return subprocess.run(cmd, shell=True, cwd=cwd, stdout=f, stderr=subprocess.STDOUT).returncode


def build_shim(cuda: bool):
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    out = os.path.join(HERE, "build", "libb200mpi.so")
    src = os.path.join(HERE, "mpishim", "mpishim.c")
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-o", out, src, "-lpthread"]
    if cuda and os.path.exists("/usr/local/cuda/include/cuda_runtime.h"):
        cmd = ["gcc", "-O2", "-fPIC", "-shared", "-DB200_MPISHIM_CUDA", "-I/usr/local/cuda/include", "-o", out,
               src, "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64", "-lcudart", "-lpthread", "-ldl"]
    subprocess.check_call(cmd)
    return out


def main():
    if "--check" in sys.argv:
        for v in ("v1_serial", "v2_mpi_only/2.1_broadcast_all", "v2_mpi_only/2.2_scatter_halo",
                  "v3_cuda_only", "v4_mpi_cuda"):
            t = os.path.join(DST, "src", "final_project", v, "template")
            print(f"{v}: {'built' if os.path.exists(t) else 'MISSING'}")
        return 0
    if not os.path.isdir(REF):
        print(f"reference not found at {REF}", file=sys.stderr)
        return 1
    os.makedirs(DST, exist_ok=True)
    log = os.path.join(DST, "build.log")
    open(log, "w").close()
    src = os.path.join(DST, "src")
    for sub in ("final_project/v1_serial", "final_project/v2_mpi_only", "final_project/v3_cuda_only",
                "final_project/v4_mpi_cuda", "homeworks/hw1"):
        d = os.path.join(src, sub)
        if os.path.exists(d):
            shutil.rmtree(d)
        shutil.copytree(os.path.join(REF, sub), d, ignore=shutil.ignore_patterns("template", "*.o", "logs*"))
    build_shim(cuda=True)
    env_path = f'PATH="{HERE}/bin:$PATH"'
    results = {}
    fp = os.path.join(src, "final_project")
    results["v1"] = sh("make clean; make", os.path.join(fp, "v1_serial"), log)
    for v in ("2.1_broadcast_all", "2.2_scatter_halo"):
        results[f"v{v[:3]}"] = sh(f"{env_path} make clean; {env_path} make", os.path.join(fp, "v2_mpi_only", v), log)
    results["v3"] = sh(f'make clean; make GPU_ARCH_FLAGS="{SM100}"', os.path.join(fp, "v3_cuda_only"), log)
    results["v4"] = sh(f'{env_path} make clean; {env_path} make GPU_ARCH_FLAGS="{SM100}"',
                       os.path.join(fp, "v4_mpi_cuda"), log)
    hw1 = os.path.join(src, "homeworks", "hw1")
    results["hw1"] = sh(f"{env_path} mpicc -O2 -o template src/template.c", hw1, log)
    for k, rc in results.items():
        print(f"{k}: {'ok' if rc == 0 else f'FAILED rc={rc}'}")
    print(f"log: {log}")
    return 0 if all(rc == 0 for rc in results.values()) else 2


if __name__ == "__main__":
    sys.exit(main())
