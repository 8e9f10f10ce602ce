"""Host-side placement for the serving path: keep a worker process (and the pinned buffers it allocates) on the NUMA node
its GPU hangs off.

``predict_host`` moves 245 KB per clip over PCIe; on an 8-GPU node with two sockets a rank whose pinned staging memory sits
on the other socket pays the inter-socket link on every copy (measured: 8 ranks reached 0.76 M clips/s end to end, fewer
than the 0.87 M of 4 ranks, without this binding).  Linux places pages on the node of the thread that first touches them, so
binding the process to the GPU's node *before* it allocates pinned memory is enough — no libnuma needed.  Everything here
is best-effort: without sysfs, on a single-node host, or where sysfs reports numa_node = -1 (virtualised boxes) it does nothing.
"""
from __future__ import annotations

import os
from typing import Optional

import torch


def _parse_cpulist(text: str) -> set:
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def pci_address(device: int) -> Optional[str]:
    """``dddd:bb:dd.0`` of a CUDA device, as sysfs names it."""
    try:
        p = torch.cuda.get_device_properties(device)
        return f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    except Exception:
        return None


def numa_node_of(device: int, sysfs: str = "/sys") -> Optional[int]:
    addr = pci_address(device)
    if addr is None:
        return None
    try:
        with open(f"{sysfs}/bus/pci/devices/{addr}/numa_node") as f:
            node = int(f.read().strip())
    except (OSError, ValueError):
        return None
    return node if node >= 0 else None


def bind_to_numa_node(node: int, sysfs: str = "/sys") -> bool:
    """Restrict this process to the CPUs of NUMA node ``node`` (intersected with the CPUs it may already use)."""
    try:
        with open(f"{sysfs}/devices/system/node/node{node}/cpulist") as f:
            cpus = _parse_cpulist(f.read())
        allowed = cpus & os.sched_getaffinity(0)
        if not allowed:
            return False
        os.sched_setaffinity(0, allowed)
    except (OSError, ValueError, AttributeError):
        return False
    return True


def bind_to_device_numa_node(device: int, sysfs: str = "/sys") -> Optional[int]:
    """Call once per worker process, before allocating pinned memory.  Returns the node, or None when nothing changed."""
    node = numa_node_of(device, sysfs)
    if node is None or not bind_to_numa_node(node, sysfs):
        return None
    return node
