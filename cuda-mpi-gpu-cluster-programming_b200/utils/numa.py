"""Host-side placement: run this process on the CPUs that are local to its GPU.

Pinned staging buffers are placed by first touch, so a process that happens to run on the other socket gets
its H2D/D2H traffic routed across the socket interconnect; the end-to-end rate then depends on where the
scheduler put the process (70k vs 85k images/s were both measured for the same code on different boxes).
The reference has nothing comparable (pageable synchronous cudaMemcpy, ref v3_cuda_only/src/alexnet_cuda.cu:67-71).

Two sources of the GPU's CPU set, tried in order: NVML's affinity mask, then sysfs
(``/sys/bus/pci/devices/<bdf>/numa_node`` -> ``/sys/devices/system/node/node<k>/cpulist``).  The report always says
what happened and why, so a bench line can show "nothing to narrow" instead of a bare null.
"""
from __future__ import annotations

import os
from typing import Optional, Set

import torch


def _parse_cpulist(text: str) -> Set[int]:
    out: Set[int] = set()
    for part in text.strip().split(","):
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            out.update(range(int(a), int(b) + 1))
        else:
            out.add(int(part))
    return out


def _nvml_cpus(idx: int) -> Optional[Set[int]]:
    import pynvml
    pynvml.nvmlInit()
    uuid = str(torch.cuda.get_device_properties(idx).uuid)
    try:
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
    except Exception:
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
    ncpu = max(os.cpu_count() or 1, max(os.sched_getaffinity(0)) + 1)
    words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
    cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
    return cpus or None


def _sysfs_cpus(idx: int) -> Optional[Set[int]]:
    bdf = torch.cuda.get_device_properties(idx).pci_bus_id if hasattr(torch.cuda.get_device_properties(idx), "pci_bus_id") else None
    if not bdf:
        import pynvml
        pynvml.nvmlInit()
        bdf = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(idx)).busId
        bdf = bdf.decode() if isinstance(bdf, bytes) else bdf
    bdf = bdf.lower()
    if len(bdf.split(":")[0]) == 8:                   # NVML prints an 8-digit domain, sysfs uses 4
        bdf = bdf[4:]
    node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
    if node < 0:
        return None
    return _parse_cpulist(open(f"/sys/devices/system/node/node{node}/cpulist").read()) or None


def bind_host_to_gpu(device) -> dict:
    """Restrict the calling process to the CPUs local to ``device`` (intersected with the CPUs it may already use).
    Never raises.  Returns ``{"action": "bound" | "none", "reason": ..., "source": ..., "cpus_before": n, "cpus_after": m}``."""
    rep = {"action": "none", "reason": "", "source": None}
    if os.environ.get("B200_NO_NUMA_BIND") == "1":
        rep["reason"] = "disabled by B200_NO_NUMA_BIND=1"
        return rep
    if not hasattr(os, "sched_setaffinity"):
        rep["reason"] = "no sched_setaffinity on this platform"
        return rep
    try:
        dev = torch.device(device)
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
        allowed = os.sched_getaffinity(0)
        rep["cpus_before"] = len(allowed)
        local, errs = None, []
        for name, fn in (("nvml", _nvml_cpus), ("sysfs", _sysfs_cpus)):
            try:
                local = fn(idx)
            except Exception as ex:
                errs.append(f"{name}: {type(ex).__name__}: {str(ex)[:80]}")
                local = None
            if local:
                rep["source"] = name
                break
        if not local:
            rep["reason"] = "no CPU locality information for this GPU (" + ("; ".join(errs) or "single NUMA node") + ")"
            return rep
        target = local & allowed
        if target == allowed:
            rep["reason"] = "every allowed CPU is already local to the GPU: nothing to narrow"
            return rep
        if len(target) < 4:
            rep["reason"] = f"only {len(target)} allowed CPU(s) are local to the GPU: too few to drive it, left unbound"
            return rep
        os.sched_setaffinity(0, target)
        rep.update(action="bound", reason="process restricted to the GPU's local CPUs", cpus_after=len(target))
        return rep
    except Exception as ex:
        rep["reason"] = f"{type(ex).__name__}: {str(ex)[:120]}"
        return rep
