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


def _topo_cpu_affinity(index: int):
    """CPU list of GPU `index` from `nvidia-smi topo -m` (works where sysfs reports numa_node = -1)."""
    import subprocess
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        header = None
        for line in out.splitlines():
            cols = [c.strip() for c in line.split("\t") if c.strip()]
            if not cols:
                continue
            if cols[0].startswith("GPU0") and header is None and "CPU Affinity" in line:
                header = cols
            if "CPU Affinity" in line and header is None:
                header = ["name"] + cols
            if cols[0] == f"GPU{index}" and header is not None:
                pos = header.index("CPU Affinity") if "CPU Affinity" in header else None
                if pos is not None and pos < len(cols):
                    return _parse_cpulist(cols[pos])
    except Exception:
        pass
    return None


def bind_to_gpu_numa(index: int) -> Optional[int]:
    """Best effort; returns the NUMA node bound to (or -2 when bound from `nvidia-smi topo`), or None
    if nothing was changed."""
    node = gpu_numa_node(index)
    if node is None:
        cpus = _topo_cpu_affinity(index)
        if cpus:
            try:
                allowed = set(os.sched_getaffinity(0))
                cpus = [c for c in cpus if c in allowed]
                if cpus and len(cpus) < len(allowed):
                    os.sched_setaffinity(0, cpus)
                    return -2
            except Exception:
                pass
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
