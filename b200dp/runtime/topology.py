"""Topology discovery: GPUs, the NVLink/NVSwitch peer-access matrix, NUMA nodes and GPU<->NUMA affinity.

Open MPI and Horovod learn this implicitly at ``MPI_Init`` (``horovod/tensorflow_mnist.py:90``; hwloc inside
Open MPI, built at ``horovod/Dockerfile:51-61``).  Here it is explicit: ``libb200dp`` reports the CUDA side
(``b200_topology_json``), ``/sys`` gives the host side, and the launcher uses it for ``-bind-to numa``.
"""
from __future__ import annotations

import os
from typing import Any, Dict, List


def _read(path: str) -> str:
    try:
        return open(path).read().strip()
    except OSError:
        return ""


def _cpulist(txt: str) -> List[int]:
    out: List[int] = []
    for part in txt.split(","):
        if "-" in part:
            a, b = part.split("-")
            out.extend(range(int(a), int(b) + 1))
        elif part:
            out.append(int(part))
    return out


def numa_nodes() -> Dict[int, List[int]]:
    base = "/sys/devices/system/node"
    nodes: Dict[int, List[int]] = {}
    if os.path.isdir(base):
        for d in sorted(os.listdir(base)):
            if d.startswith("node") and d[4:].isdigit():
                nodes[int(d[4:])] = _cpulist(_read(os.path.join(base, d, "cpulist")))
    return nodes


def gpu_numa_node(pci_bus_id: str) -> int:
    v = _read(f"/sys/bus/pci/devices/{pci_bus_id.lower()}/numa_node")
    try:
        return int(v)
    except ValueError:
        return -1


def discover() -> Dict[str, Any]:
    """``{"gpus": [...], "peer_access": [[...]], "numa": {...}, "all_to_all": bool, "nvls": bool}``"""
    info: Dict[str, Any] = {"gpus": [], "peer_access": [], "numa": {str(k): v for k, v in numa_nodes().items()}}
    try:
        import torch
        if torch.cuda.is_available():
            from ..ops import kernels as K
            info.update(K.topology())
            for g in info["gpus"]:
                g["numa_node"] = gpu_numa_node(g.get("pci", ""))
            pa = info.get("peer_access", [])
            info["all_to_all"] = bool(pa) and all(all(v == 1 for v in row) for row in pa)
            info["nvls"] = bool(K.lib().b200_nvls_supported())
    except Exception as e:  # noqa: BLE001
        info["error"] = str(e)
    info["host_cpus"] = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    return info


if __name__ == "__main__":
    import json
    print(json.dumps(discover(), indent=1))
