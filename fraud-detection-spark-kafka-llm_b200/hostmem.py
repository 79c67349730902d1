"""Host-side placement for the host-buffer path (``fd_transform_host``): on a multi-socket box every GPU hangs off one
socket, and pinned buffers that live on the other socket's memory cross the inter-socket link on every copy -- with
one process per GPU all copying at once, that link (not PCIe) sets the end-to-end rate.  ``numa_local(device)`` runs
the allocation of the pinned buffers on CPUs of the GPU's own NUMA node, so first touch puts the pages there."""
from __future__ import annotations

import contextlib
import os
from typing import Dict, List, Optional


def _parse_cpulist(text: str) -> List[int]:
    cpus: List[int] = []
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.extend(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_numa_node(device: int) -> Optional[int]:
    """NUMA node of CUDA device ``device`` from sysfs (None when the box has one node or does not say)."""
    import torch

    p = torch.cuda.get_device_properties(device)
    bdf = "%04x:%02x:%02x.0" % (getattr(p, "pci_domain_id", 0), p.pci_bus_id, p.pci_device_id)
    try:
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            node = int(f.read().strip())
    except (OSError, ValueError):
        return None
    return node if node >= 0 else None


def node_cpus(node: int) -> List[int]:
    try:
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            return _parse_cpulist(f.read())
    except OSError:
        return []


@contextlib.contextmanager
def numa_local(device: int, report: Optional[Dict] = None):
    """Context manager: inside it the calling thread runs on the CPUs of ``device``'s NUMA node (allocate and first-touch
    the pinned buffers there); the previous affinity is restored on exit.  A box with a single node, or without the
    sysfs entries, is left alone.  ``report`` (a dict) receives what was done."""
    info = {"node": None, "bound": False}
    old = None
    try:
        node = gpu_numa_node(device)
        info["node"] = node
        if node is not None and hasattr(os, "sched_setaffinity"):
            allowed = os.sched_getaffinity(0)
            cpus = set(node_cpus(node)) & allowed
            if cpus and cpus != allowed:
                old = allowed
                os.sched_setaffinity(0, cpus)
                info["bound"] = True
                info["cpus"] = len(cpus)
    except OSError:
        pass
    if report is not None:
        report.update(info)
    try:
        yield info
    finally:
        if old is not None:
            try:
                os.sched_setaffinity(0, old)
            except OSError:
                pass
