"""Pin the calling process to the CPUs next to its GPU, so that pinned host
buffers are first-touched on the right NUMA node (matters for the end-to-end
number at 8 GPUs: eight ranks pulling from one memory controller share ~100 GB/s)."""
from __future__ import annotations

import os


def _parse_cpulist(s: str):
    cpus = set()
    for part in s.strip().split(","):
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def bind_to_gpu_numa(device_index: int) -> dict:
    """Best effort; returns what it did (for logging).  Never raises."""
    info = {"device": device_index, "bound": False}
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        bus = f"{getattr(p, 'pci_domain_id', 0):04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        info["pci"] = bus
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            node = int(f.read().strip())
        info["numa_node"] = node
        if node < 0:
            return info
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = _parse_cpulist(f.read())
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if cpus:
            os.sched_setaffinity(0, cpus)
            info["bound"] = True
            info["cpus"] = len(cpus)
    except Exception as e:  # noqa: BLE001
        info["error"] = str(e)[:80]
    return info
