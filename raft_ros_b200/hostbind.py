"""Host-side placement for the end-to-end path: run the process, and first-touch its pinned buffers, on the
NUMA node its GPU hangs off.

The correlation results of a step are large (0.88 GB at batch 8 @ 440x1024) and with one process per GPU all
of them land in host memory at once; which memory controller they land on is decided by where the pages
of the pinned buffers are first touched.  Nothing here is on the device path.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional


def _read(path: str) -> Optional[str]:
    try:
        with open(path) as f:
            return f.read().strip()
    except OSError:
        return None


def _parse_cpulist(s: str) -> List[int]:
    cpus: List[int] = []
    for part in s.split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.extend(range(int(a), int(b) + 1))
        else:
            cpus.append(int(part))
    return cpus


def gpu_numa_node(index: int) -> Optional[int]:
    """NUMA node of CUDA device `index` from sysfs (None when the platform does not say)."""
    import torch

    p = torch.cuda.get_device_properties(index)
    try:
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    except AttributeError:
        return None
    v = _read(f"/sys/bus/pci/devices/{bdf}/numa_node")
    if v is None:
        return None
    node = int(v)
    return node if node >= 0 else None


def node_cpus(node: int) -> List[int]:
    s = _read(f"/sys/devices/system/node/node{node}/cpulist")
    return _parse_cpulist(s) if s else []


def bind_near_gpu(index: int, rank_on_node: int = 0, ranks_on_node: int = 1) -> Dict:
    """Restrict this process to the CPUs of the GPU's NUMA node (a disjoint slice of them per rank when several
    ranks share the node), so that pages it touches from now on are allocated there.  Returns what was done."""
    allowed = sorted(os.sched_getaffinity(0))
    info: Dict = {"gpu": index, "numa_node": None, "cpus_before": len(allowed), "bound": False}
    node = gpu_numa_node(index)
    info["numa_node"] = node
    if node is None:
        return info
    cpus = [c for c in node_cpus(node) if c in allowed]
    if not cpus:
        return info
    if ranks_on_node > 1 and len(cpus) >= ranks_on_node:
        per = len(cpus) // ranks_on_node
        cpus = cpus[rank_on_node * per:(rank_on_node + 1) * per]
    try:
        os.sched_setaffinity(0, cpus)
        info.update(bound=True, cpus_after=len(cpus))
    except OSError as e:  # containers may forbid it
        info["error"] = str(e)
    return info


def topology() -> Dict:
    """{numa nodes: cpu counts, gpu -> node}: printed next to the end-to-end numbers."""
    import torch

    nodes = {}
    n = 0
    while os.path.isdir(f"/sys/devices/system/node/node{n}"):
        nodes[n] = len(node_cpus(n))
        n += 1
    return {"numa_cpus": nodes, "gpu_numa": {i: gpu_numa_node(i) for i in range(torch.cuda.device_count())},
            "affinity": len(os.sched_getaffinity(0))}


class near_gpu:
    """`with near_gpu(index, rank, ranks): buf = torch.empty(...).pin_memory()` -- the calling thread is bound to the
    GPU's NUMA node while the block runs (so the pinned pages are allocated there) and released afterwards (so that
    later CPU work, e.g. a multi-threaded baseline, sees all cores again)."""

    def __init__(self, index: int, rank_on_node: int = 0, ranks_on_node: int = 1):
        self.args = (index, rank_on_node, ranks_on_node)
        self.info: Dict = {}
        self._saved = None

    def __enter__(self):
        self._saved = os.sched_getaffinity(0)
        self.info = bind_near_gpu(*self.args)
        return self

    def __exit__(self, *exc):
        try:
            os.sched_setaffinity(0, self._saved)
        except OSError:
            pass
        return False
