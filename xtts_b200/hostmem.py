"""Host-side placement for the end-to-end path: keep a rank's threads and its pinned staging buffers on the NUMA
node its GPU hangs off.

One process per GPU moves ~74 MB per 64 x 10 s step over PCIe (fp32 waveform in, log-mel out).  A pinned buffer that
was first touched on the other socket crosses the inter-socket link on every copy, and several ranks doing so share
that link: on an 8-GPU box the host-buffer throughput then stops scaling long before the GPUs do.  Pure host logic
(sysfs + sched_setaffinity), no CUDA calls beyond reading the device's PCI address.
"""
from __future__ import annotations

import os
from typing import Optional, Set


def _parse_cpulist(text: str) -> Set[int]:
    cpus: Set[int] = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_numa_node(device_index: int, sysfs: str = "/sys") -> Optional[int]:
    """NUMA node of a CUDA device from its PCI address, or None when the platform does not say (single-socket
    hosts and most VMs report -1)."""
    import torch
    p = torch.cuda.get_device_properties(device_index)
    addr = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    try:
        node = int(open(f"{sysfs}/bus/pci/devices/{addr}/numa_node").read().strip())
    except (OSError, ValueError):
        return None
    return node if node >= 0 else None


def bind_to_gpu_numa(device_index: int, sysfs: str = "/sys") -> dict:
    """Restrict this process to the CPUs of the GPU's NUMA node (so that pinned buffers allocated afterwards are
    first-touched there).  Returns what was done: {"node": n | None, "cpus": count}."""
    node = gpu_numa_node(device_index, sysfs)
    if node is None:
        return {"node": None, "cpus": len(os.sched_getaffinity(0))}
    try:
        cpus = _parse_cpulist(open(f"{sysfs}/devices/system/node/node{node}/cpulist").read())
        allowed = cpus & os.sched_getaffinity(0)
        if allowed:
            os.sched_setaffinity(0, allowed)
        return {"node": node, "cpus": len(os.sched_getaffinity(0))}
    except OSError:
        return {"node": node, "cpus": len(os.sched_getaffinity(0))}


def unbind(cpus) -> None:
    """Give every thread of this process the CPU set `cpus` again (e.g. the set saved before ``bind_to_gpu_numa``):
    worker pools created while bound inherited the narrow mask."""
    for tid in os.listdir("/proc/self/task"):
        try:
            os.sched_setaffinity(int(tid), cpus)
        except (OSError, ValueError):
            pass
