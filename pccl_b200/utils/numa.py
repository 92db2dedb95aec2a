"""Keeps a peer's host threads and page-locked buffers next to its GPU.

On a multi-socket 8-GPU box every GPU hangs off one CPU socket.  A process that was started without any binding lands
on an arbitrary core; its pinned staging buffers are then allocated (first touch) on that core's memory node, and when
eight such processes pull 600 MB each through one memory controller the host-to-device copies run at a fraction of PCIe
speed (measured in round 1: 22 GB/s per GPU with eight peers against 55 GB/s alone).  NCCL binds its proxy threads the
same way.  The reference has nothing of the kind (its peers are plain processes, docs/md/05-ImplementationNotes)."""
from __future__ import annotations

import os
from pathlib import Path


def _parse_cpulist(text: str) -> set[int]:
    cpus: set[int] = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def device_numa_node(device_index: int) -> int | None:
    """NUMA node of a CUDA device, from sysfs; None when it cannot be told (single node, container without sysfs)."""
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        bus = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        node = int(Path(f"/sys/bus/pci/devices/{bus}/numa_node").read_text().strip())
        return node if node >= 0 else None
    except Exception:
        return None


def _affinity_from_nvidia_smi(device_index: int) -> set[int]:
    """CPU affinity of a GPU as `nvidia-smi topo -m` reports it (containers often hide the sysfs NUMA node)."""
    import re
    import subprocess
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=20).stdout
    except Exception:
        return set()
    for line in out.splitlines():
        cells = [c for c in re.split(r"\t+|\s{2,}", re.sub(r"\x1b\[[0-9;]*m", "", line).strip()) if c]
        if not cells or cells[0] != f"GPU{device_index}":
            continue
        for c in cells[1:]:
            if re.fullmatch(r"\d+(-\d+)?(,\d+(-\d+)?)*", c) and ("-" in c or "," in c):
                return _parse_cpulist(c)
    return set()


def bind_to_device(device_index: int) -> dict:
    """Restricts the calling process to the CPUs of the GPU's memory node (so later allocations, pinned ones included,
    are local to it).  Does nothing, and says why, when the topology is unknown.  Returns a small report."""
    node = device_numa_node(device_index)
    if node is None:
        cpus = _affinity_from_nvidia_smi(device_index)
        try:
            target = cpus & os.sched_getaffinity(0)
            if target and len(target) < len(os.sched_getaffinity(0)):
                os.sched_setaffinity(0, target)
                return {"bound": True, "source": "nvidia-smi topo", "cpus": len(target)}
        except Exception as e:  # noqa: BLE001
            return {"bound": False, "reason": str(e)}
        return {"bound": False, "reason": "no NUMA information for the device"}
    try:
        cpus = _parse_cpulist(Path(f"/sys/devices/system/node/node{node}/cpulist").read_text())
        allowed = os.sched_getaffinity(0)
        target = cpus & allowed
        if not target:
            return {"bound": False, "node": node, "reason": "none of the node's CPUs is available to this process"}
        os.sched_setaffinity(0, target)
        return {"bound": True, "node": node, "cpus": len(target)}
    except Exception as e:  # noqa: BLE001
        return {"bound": False, "node": node, "reason": str(e)}
