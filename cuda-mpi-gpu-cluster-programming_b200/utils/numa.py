"""Host-side placement: run this process on the CPUs that are local to its GPU.

Pinned staging buffers are placed by first touch, so a process that happens to run on the other socket gets
its H2D/D2H traffic routed across the socket interconnect; the end-to-end rate then depends on where the
scheduler put the process (70k vs 85k images/s were both measured for the same code on different boxes).
The reference has nothing comparable (pageable synchronous cudaMemcpy, ref v3_cuda_only/src/alexnet_cuda.cu:67-71).
"""
from __future__ import annotations

import os
from typing import Optional

import torch


def bind_host_to_gpu(device) -> Optional[dict]:
    """Restrict the calling process to the CPUs NVML reports as local to ``device`` (intersected with the CPUs
    the process may already use).  Returns a small report, or None when nothing was changed (no NVML, no
    affinity information, nothing to narrow, B200_NO_NUMA_BIND=1).  Never raises."""
    if os.environ.get("B200_NO_NUMA_BIND") == "1" or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        import pynvml
        dev = torch.device(device)
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(idx).uuid)
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        allowed = os.sched_getaffinity(0)
        ncpu = max(os.cpu_count() or 1, max(allowed) + 1)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        local = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        target = local & allowed
        if len(target) < 4 or target == allowed:       # nothing to narrow, or too few cores left to drive the GPU
            return None
        os.sched_setaffinity(0, target)
        return {"cpus_before": len(allowed), "cpus_after": len(target)}
    except Exception:
        return None
