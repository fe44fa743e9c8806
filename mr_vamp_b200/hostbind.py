"""Host-side placement for the host-buffer entry points (one process per GPU).

The end-to-end path copies 28 B per configuration over PCIe; with eight ranks on one two-socket box every rank's
staging memory and copy-issuing thread should sit on the NUMA node its GPU hangs off, otherwise all copies cross one
socket's memory controller and inter-socket link (round-1 scaling run: e2e efficiency 0.42 at 8 GPUs with every rank on
node 0).  `bind_to_gpu` pins the calling process to the GPU's local CPUs BEFORE pinned buffers are allocated, so that
their pages are first-touched on that node.  Nothing here computes anything of the path.
"""
from __future__ import annotations

import ctypes as C
import os


def _read(path: str):
    try:
        with open(path) as fh:
            return fh.read().strip()
    except OSError:
        return None


def _parse_cpulist(text: str) -> set[int]:
    cpus: set[int] = set()
    for part in (text or "").split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def describe(device: int) -> dict:
    """PCI id, NUMA node and local CPUs of a CUDA device (from sysfs); what the process is currently allowed to run on."""
    from . import api

    buf = C.create_string_buffer(32)
    api._check(api.lib().vb_device_pci_bus_id(device, buf, 32))
    bus = buf.value.decode().lower()
    base = f"/sys/bus/pci/devices/{bus}"
    node = _read(f"{base}/numa_node")
    cpus = _parse_cpulist(_read(f"{base}/local_cpulist"))
    return {"pci": bus, "numa_node": int(node) if node not in (None, "") else None, "local_cpus": len(cpus),
            "allowed_cpus": len(os.sched_getaffinity(0)), "_cpus": cpus}


def bind_to_gpu(device: int, share: tuple[int, int] | None = None) -> dict:
    """Pin this process to the CPUs local to `device`'s NUMA node (all of them, or the share[0]-th of share[1] equal
    slices when several ranks share a node).  Returns describe() with what was applied; a box without NUMA information
    is left untouched."""
    info = describe(device)
    cpus = sorted(info.pop("_cpus") & os.sched_getaffinity(0))
    if cpus and share:
        k, n = share
        per = max(1, len(cpus) // n)
        cpus = cpus[k * per:(k + 1) * per] or cpus
    if cpus:
        os.sched_setaffinity(0, cpus)
        info["bound_cpus"] = len(cpus)
    else:
        info["bound_cpus"] = 0
    return info
