"""Keep a replica's host threads — and with them the pinned staging buffers they allocate — on the CPUs next to its GPU.

On a multi-GPU box every replica streams its images through host memory (pinned upload / download buffers).  Linux places
pinned pages on the NUMA node of the allocating thread, so a replica that runs on the far socket sends every byte across
the socket interconnect before it reaches its GPU's PCIe root.  `bind_host_to_gpu` asks NVML which CPUs are local to the
device and restricts the process to those it is allowed to use (the reference leaves placement to Ray's `num_cpus`, ref:
apps/model-runner/runtime_deployment.py:33-36).  Call it before allocating pinned memory.  BE_B200_NO_NUMA_BIND=1 turns
it into a no-op.
"""
from __future__ import annotations

import os
from typing import Dict, Optional


def bind_host_to_gpu(device_index: int) -> Dict[str, object]:
    """→ {"bound": bool, "cpus": number of CPUs the process may use now, "reason": why not, if not}."""
    info: Dict[str, object] = {"bound": False, "cpus": len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else None}
    if os.environ.get("BE_B200_NO_NUMA_BIND") == "1":
        info["reason"] = "disabled by BE_B200_NO_NUMA_BIND"
        return info
    if not hasattr(os, "sched_setaffinity"):
        info["reason"] = "no sched_setaffinity on this platform"
        return info
    try:
        import pynvml
        import torch

        pynvml.nvmlInit()
        try:
            handle: Optional[object] = None
            props = torch.cuda.get_device_properties(device_index)
            if all(hasattr(props, a) for a in ("pci_domain_id", "pci_bus_id", "pci_device_id")):
                bus = f"{props.pci_domain_id:08x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
                try:
                    handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
                except Exception:
                    handle = None
            if handle is None:
                handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            n_cpu = os.cpu_count() or 1
            words = pynvml.nvmlDeviceGetCpuAffinity(handle, (n_cpu + 63) // 64)
            local = {64 * w + b for w, mask in enumerate(words) for b in range(64) if (int(mask) >> b) & 1}
        finally:
            pynvml.nvmlShutdown()
        allowed = os.sched_getaffinity(0)
        want = local & allowed
        if not want:
            info["reason"] = "none of the GPU's local CPUs is available to this process"
            return info
        if want == allowed:
            info["reason"] = "already on the GPU's local CPUs"
            info["bound"] = True
            return info
        os.sched_setaffinity(0, want)
        info.update(bound=True, cpus=len(want), cpus_before=len(allowed))
    except Exception as e:   # NVML missing, permissions, exotic topologies: placement stays as it was
        info["reason"] = f"{type(e).__name__}: {e}"
    return info
