"""Bind a one-process-per-GPU rank to the CPUs (and therefore the memory) of its GPU's NUMA node.

Pinned host buffers are allocated on the node of the allocating thread; a source that copies 64 MiB per round from pinned
memory wants that memory next to its GPU's PCIe root.  Best effort: any missing sysfs entry or a refused affinity call leaves
the process as it was."""
from __future__ import annotations

import os
from pathlib import Path


def _parse_cpulist(text: str) -> set[int]:
    cpus: set[int] = set()
    for part in text.strip().split(","):
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def gpu_numa_node(device: int) -> int | None:
    try:
        import torch

        p = torch.cuda.get_device_properties(device)
        bus = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        node = int((Path("/sys/bus/pci/devices") / bus / "numa_node").read_text())
        return node if node >= 0 else None
    except Exception:
        return None


def bind_to_gpu_numa(device: int) -> dict:
    """Returns {"node": n, "cpus": k} when the binding was applied, {} otherwise."""
    try:
        node = gpu_numa_node(device)
        if node is None:
            return {}
        cpus = _parse_cpulist((Path("/sys/devices/system/node") / f"node{node}" / "cpulist").read_text())
        allowed = os.sched_getaffinity(0)
        target = cpus & allowed
        if not target or target == allowed:
            return {}
        os.sched_setaffinity(0, target)
        return {"node": node, "cpus": len(target)}
    except Exception:
        return {}
