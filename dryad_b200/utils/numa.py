"""NUMA placement of a rank's host side: pin the process to the CPUs that are local to its GPU so
that pinned staging buffers (first-touch) and the threads that fill them sit on the GPU's socket.
On a two-socket 8-GPU box an H2D copy from the remote socket crosses the inter-socket link and all
eight ranks contend for it; the reference's analogue is the scheduler's machine/rack affinity for
vertex inputs (GraphManager/kernel/DrCluster.cpp, affinity weights)."""
from __future__ import annotations

import os
from typing import List, Optional


def gpu_local_cpus(device_index: int) -> Optional[List[int]]:
    """CPUs NVML reports as ideal for this GPU, or None when NVML is unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        ncpu = os.cpu_count() or 1
        words = (ncpu + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [w * 64 + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1]
        return [c for c in cpus if c < ncpu] or None
    except Exception:  # noqa: BLE001
        return None


def bind_to_gpu(device_index: int) -> Optional[List[int]]:
    """Restrict the calling process to the GPU-local CPUs (intersected with what it may already
    use).  Returns the CPU list applied, or None when nothing was changed.  DRYAD_NUMA_BIND=0
    disables it."""
    if os.environ.get("DRYAD_NUMA_BIND", "1") == "0" or not hasattr(os, "sched_setaffinity"):
        return None
    cpus = gpu_local_cpus(device_index)
    if not cpus:
        return None
    allowed = os.sched_getaffinity(0)
    use = sorted(set(cpus) & allowed)
    if not use or len(use) == len(allowed):
        return None
    try:
        os.sched_setaffinity(0, use)
    except OSError:
        return None
    return use
