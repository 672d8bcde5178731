"""`bench.py --impl reference`: the UNMODIFIED reference, run through its own public interface —
the `template` executable (`mpirun -np N ./template` for V4) — on the same box.

The reference has no batch dimension and parses no arguments: one run = one
227x227x3 image, constant init, and its own wall-clock around the whole forward (context creation,
cudaMalloc/cudaFree, pageable H2D/D2H and, for V4, the MPI scatter/halo/gather included,
ref v4_mpi_cuda/src/main_mpi_cuda.cpp:151-159).  images/s = 1 / (its printed time); a batch of 128
is 128 such runs.  Each step below is one run; W warm-up runs are discarded.

Build: baseline/build_reference.py (copies the sources to baseline/_ref/src verbatim; nvcc
-gencode arch=compute_100,code=sm_100; MPI = baseline/mpishim because the image has no MPI).
One GPU per rank through CUDA_VISIBLE_DEVICES (the reference never calls cudaSetDevice).
"""
import json
import os
import re
import statistics
import subprocess
import sys

from bench_contract import METRIC, UNIT, HIGHER_IS_BETTER, SCALING, MODEL

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
V4_DIR = os.path.join(HERE, "_ref", "src", "final_project", "v4_mpi_cuda")
RE_T = re.compile(r"AlexNet MPI\+CUDA Forward Pass completed in ([0-9.eE+-]+) ms")


def unavailable(why: str):
    print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)


def main(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:            # the reference is launched by its own mpirun line from rank 0 only
        return
    N = int(os.environ.get("WORLD_SIZE", a.gpus))
    exe = os.path.join(V4_DIR, "template")
    if not os.path.exists(exe):
        try:
            subprocess.run([sys.executable, os.path.join(HERE, "build_reference.py")], check=False,
                           capture_output=True, timeout=900)
        except Exception:
            pass
    if not os.path.exists(exe):
        return unavailable("baseline/_ref V4 binary is not built (baseline/build_reference.py failed; "
                           "pip cannot install the reference: no setup.py/pyproject)")
    env = dict(os.environ, B200_MPIRUN_BIND_GPUS="1", B200_MPIRUN_NUM_GPUS=str(N))
    env["LD_LIBRARY_PATH"] = "/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")   # -cudart=shared
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_PORT", "MASTER_ADDR", "OMPI_COMM_WORLD_RANK",
              "OMPI_COMM_WORLD_SIZE", "CUDA_VISIBLE_DEVICES", "TORCHELASTIC_RUN_ID"):
        env.pop(k, None)
    cmd = [sys.executable, os.path.join(ROOT, "bin", "mpirun"), "--oversubscribe", "-np", str(N), "./template"]
    times, shape = [], None
    for i in range(a.warmup + a.steps):
        try:
            p = subprocess.run(cmd, cwd=V4_DIR, env=env, capture_output=True, text=True, timeout=300)
        except subprocess.TimeoutExpired:
            return unavailable("reference V4 run timed out")
        m = RE_T.search(p.stdout)
        if p.returncode != 0 or not m:
            return unavailable(f"reference V4 run failed rc={p.returncode}: {(p.stderr or p.stdout)[-200:]!r}")
        s = re.search(r"^Final Output Shape: (\S+)", p.stdout, re.M)
        shape = s.group(1) if s else None
        if i >= a.warmup:
            times.append(float(m.group(1)))
    ms = statistics.mean(times)
    val = 1000.0 / ms
    print(json.dumps({
        "impl": "reference",
        "metric": METRIC,
        "value": round(val, 3), "unit": UNIT, "n_gpus": N, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": round(ms, 3), "higher_is_better": HIGHER_IS_BETTER, "scaling": SCALING, "vs_baseline": None,
        "dtype": "fp32", "data": "synthetic (constant init 1.0 / 0.01 / 0.0, as the reference hard-codes)",
        "config": {"model": MODEL, "implementation": "reference V4 MPI+CUDA, unmodified sources, sm_100 rebuild",
                   "timed_region": "the reference's own wall-clock line: one 227x227x3 image per run, context creation, "
                                   "cudaMalloc/cudaFree, pageable copies and MPI traffic included (it has no batch and no warm loop)",
                   "global_batch": 1, "input": "227x227x3", "parallelism": f"mpirun -np {N} (row scatter + host halo)",
                   "output_shape_printed": shape,
                   "note": "reference output rows shrink with np (13/8/4 at np 1/2/4): time only, not a correctness target"},
        "e2e": {"value": round(val, 3), "unit": UNIT, "h2d_bytes_per_step": 3216748,
                "d2h_bytes_per_step": 173056, "note": "the reference's timed region already contains its copies"},
        "gpu_launches": 7 * N, "median_ms": round(statistics.median(times), 3),
    }), flush=True)
