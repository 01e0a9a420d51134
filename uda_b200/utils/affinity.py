"""Bind a rank to the CPUs (and thereby, through first-touch, the host memory) next to its GPU.

The end-to-end path moves every step's map outputs host -> device and the merged stream device -> host through pinned
buffers. With one process per GPU on a multi-socket host, pinned memory that lands on the other socket crosses the
inter-socket link on every copy and the ranks share that link. The reference has the same concern for its registered
RDMA buffers (``mapred.rdma.mem.use.contig.pages``, src/DataNet/RDMAComm.cc:117, picks the allocation policy); here
the policy is: run on the GPU's local CPUs before the pinned buffers are allocated.
"""
from __future__ import annotations

import os
import subprocess
from typing import Optional, Set


def _parse_cpulist(s: str) -> Set[int]:
    out: Set[int] = set()
    for part in s.strip().split(","):
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            out.update(range(int(a), int(b) + 1))
        else:
            out.add(int(part))
    return out


def gpu_pci_address(index: int) -> Optional[str]:
    try:
        import torch

        p = torch.cuda.get_device_properties(index)
        if all(hasattr(p, a) for a in ("pci_domain_id", "pci_bus_id", "pci_device_id")):
            return f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    except Exception:
        pass
    try:
        q = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(index)],
                           capture_output=True, text=True, timeout=20)
        s = q.stdout.strip().lower()        # 00000000:1b:00.0
        if s.count(":") == 2:
            dom, bus, rest = s.split(":")
            return f"{dom[-4:]}:{bus}:{rest}"
    except Exception:
        pass
    return None


def bind_to_gpu_numa(index: int) -> dict:
    """Restricts the calling process to the CPUs local to GPU ``index``. Returns what was done (for the bench log);
    never raises: on any problem the affinity is left alone."""
    info = {"gpu": index, "bound": False}
    try:
        addr = gpu_pci_address(index)
        info["pci"] = addr
        if not addr:
            return info
        base = f"/sys/bus/pci/devices/{addr}"
        with open(f"{base}/local_cpulist") as f:
            local = _parse_cpulist(f.read())
        try:
            with open(f"{base}/numa_node") as f:
                info["numa_node"] = int(f.read().strip())
        except (OSError, ValueError):
            pass
        allowed = os.sched_getaffinity(0)
        cpus = local & allowed
        info["local_cpus"], info["allowed_cpus"] = len(local), len(allowed)
        if not cpus or cpus == allowed:
            # nothing to narrow: the GPU's local CPU list is the whole allowed set (a single-NUMA-node host or VM, where
            # numa_node reads -1), or the cpuset excludes the local CPUs
            info["reason"] = "local CPU list equals the allowed set (single NUMA node)" if cpus else "cpuset excludes the GPU's local CPUs"
            return info
        os.sched_setaffinity(0, cpus)
        info["bound"], info["cpus"] = True, len(cpus)
    except Exception as e:  # noqa: BLE001 - affinity is an optimisation
        info["error"] = f"{type(e).__name__}: {e}"
    return info
