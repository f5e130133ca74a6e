"""Pin the calling process to the CPUs next to its GPU, so that pinned host
buffers are first-touched on the right NUMA node (matters for the end-to-end
number at 8 GPUs: eight ranks pulling from one memory controller share its bandwidth).

The affinity comes from sysfs when the PCI device reports a NUMA node, and from NVML
(``nvmlDeviceGetCpuAffinity``, what ``nvidia-smi topo -m`` prints) when sysfs says -1,
as it does inside many containers.  Call it BEFORE allocating pinned memory.
"""
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


def _sysfs_cpus(bus: str, info: dict):
    with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
        node = int(f.read().strip())
    info["numa_node"] = node
    if node < 0:
        return set()
    with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
        return _parse_cpulist(f.read())


def _nvml_cpus(bus: str, info: dict):
    """CPU set NVML reports as closest to the GPU with this PCI bus id."""
    import pynvml
    pynvml.nvmlInit()
    try:
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        try:
            info["nvml_numa_node"] = int(pynvml.nvmlDeviceGetNumaNodeId(h))
        except Exception:  # noqa: BLE001 - older NVML
            pass
        return cpus
    finally:
        pynvml.nvmlShutdown()


def bind_to_gpu_numa(device_index: int) -> dict:
    """Best effort; returns what it did (for logging).  Never raises."""
    info = {"device": device_index, "bound": False}
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        bus = f"{getattr(p, 'pci_domain_id', 0):04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        info["pci"] = bus
        cpus, source = set(), None
        try:
            cpus, source = _sysfs_cpus(bus, info), "sysfs"
        except OSError:
            info["numa_node"] = None
        if not cpus:
            try:
                cpus, source = _nvml_cpus(bus, info), "nvml"
            except Exception as e:  # noqa: BLE001
                info["nvml_error"] = str(e)[:80]
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        # an affinity that spans every allowed CPU says nothing (single-socket box or a
        # container that hides the topology): leave the process alone
        if cpus and cpus != allowed:
            os.sched_setaffinity(0, cpus)
            info["bound"] = True
        info["cpus"] = len(cpus)
        info["allowed_cpus"] = len(allowed)
        info["source"] = source
    except Exception as e:  # noqa: BLE001
        info["error"] = str(e)[:80]
    return info
