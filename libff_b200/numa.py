"""Host-side placement for one-process-per-GPU launches: run this process (and therefore the
first-touch placement of its pinned staging and source buffers, and the engine's copy threads) on
the CPUs of the NUMA node its GPU hangs off.  Eight ranks that each pull 2 GiB per step through a
host whose memory all sits on one node share that node's memory controllers and the inter-socket
link; placed next to their GPUs they only share their own socket.

No libnuma in the image: the topology is read from sysfs and applied with sched_setaffinity, which
also steers first-touch page placement.  Everything here is best effort: on a single-node host, in
a cpuset that does not reach the GPU's node, or without sysfs it changes nothing and says why.
"""
import os
from typing import Optional


def _parse_cpulist(text: str) -> set:
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        if "-" in part:
            lo, hi = part.split("-")
            cpus.update(range(int(lo), int(hi) + 1))
        else:
            cpus.add(int(part))
    return cpus


def gpu_numa_node(pci_bus_id: str) -> Optional[int]:
    """NUMA node of a PCI device ("0000:1b:00.0"), None when the kernel does not say"""
    for name in (pci_bus_id.lower(), pci_bus_id.lower()[-12:]):
        path = f"/sys/bus/pci/devices/{name}/numa_node"
        try:
            node = int(open(path).read().strip())
            return node if node >= 0 else None
        except (OSError, ValueError):
            continue
    return None


def node_cpus(node: int) -> set:
    try:
        return _parse_cpulist(open(f"/sys/devices/system/node/node{node}/cpulist").read())
    except OSError:
        return set()


def online_nodes() -> list:
    try:
        return sorted(_parse_cpulist(open("/sys/devices/system/node/online").read()))
    except OSError:
        return []


def bind_to_gpu_node(device_index: int, min_cpus: int = 2) -> dict:
    """Restrict this process to the CPUs of the NUMA node of CUDA device `device_index` (those of
    them the process may use).  Returns what was found and done, for the bench line."""
    info = {"device": device_index, "bound": False}
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        bus = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    except Exception as e:  # noqa: BLE001 -- report, never fail a measurement over placement
        info["why"] = f"no PCI id: {e}"
        return info
    info["pci"] = bus
    nodes = online_nodes()
    info["nodes_online"] = len(nodes)
    node = gpu_numa_node(bus)
    info["node"] = node
    if node is None or len(nodes) < 2:
        info["why"] = "single NUMA node or node unknown"
        return info
    allowed = os.sched_getaffinity(0)
    want = node_cpus(node) & allowed
    info["cpus_allowed"] = len(allowed)
    info["cpus_on_node"] = len(want)
    if len(want) < min_cpus:
        info["why"] = "the process's cpuset does not reach the GPU's node"
        return info
    os.sched_setaffinity(0, want)
    info["bound"] = True
    return info
