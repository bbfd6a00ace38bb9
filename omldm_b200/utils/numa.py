"""NUMA placement: run each rank's host threads (and first-touch its pinned ring) on the CPU socket
its GPU hangs off, so H2D copies do not cross the inter-socket link."""
from __future__ import annotations

import os
from typing import Optional


def _parse_cpulist(s: str):
    cpus = []
    for part in s.strip().split(","):
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus += list(range(int(a), int(b) + 1))
        else:
            cpus.append(int(part))
    return cpus


def gpu_numa_node(index: int) -> Optional[int]:
    try:
        import torch
        p = torch.cuda.get_device_properties(index)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            n = int(f.read().strip())
        return n if n >= 0 else None
    except Exception:
        return None


def bind_to_gpu_numa(index: int) -> Optional[int]:
    """Best effort; returns the NUMA node bound to, or None if nothing was changed."""
    node = gpu_numa_node(index)
    if node is None:
        return None
    try:
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = _parse_cpulist(f.read())
        allowed = set(os.sched_getaffinity(0))
        cpus = [c for c in cpus if c in allowed]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        pass
    return None
