"""Host placement for one-process-per-GPU runs: keep an agent's host threads and its page-locked staging buffers on the NUMA
node its GPU is attached to.

The end-to-end path moves every frame over PCIe (361 KB in, ~70 KB of keypoints / descriptors out).  On a two-socket host
with eight GPUs a rank whose pinned buffers live on the other socket pulls them over the inter-socket link, and the aggregate
rate of eight ranks sags (round 1: 0.69 scaling efficiency at 8 GPUs end to end against 0.99 device-resident).  Linux places
pages on the node of the thread that first touches them (and cudaHostAlloc touches them in the calling thread), so pinning the
PROCESS to the GPU's node before any buffer is allocated is enough; threads created later inherit the mask.

The reference runs on the CPU and has no counterpart; the one thing it shares is the shape - one Tracking thread per agent
(src/MultiAgentSystem.cc:56-66) - which `agents.MultiAgentFrontEnd` mirrors.

Everything here is best effort: a host without NUMA information in sysfs (numa_node = -1, a VM, a container that hides
/sys) is left alone and the reason is reported.
"""
import os

from ._lib import device_pci_bus_id

_SYS_PCI = "/sys/bus/pci/devices"
_SYS_NODE = "/sys/devices/system/node"


def parse_cpulist(text):
    """"0-3,8,10-11" -> {0, 1, 2, 3, 8, 10, 11} (the format of /sys/devices/system/node/nodeN/cpulist)."""
    cpus = set()
    for part in text.strip().split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            lo, hi = part.split("-", 1)
            cpus.update(range(int(lo), int(hi) + 1))
        else:
            cpus.add(int(part))
    return cpus


def node_of_pci_device(bus_id, sys_pci=_SYS_PCI):
    """NUMA node of a PCI device, or None when sysfs does not say (file missing or -1)."""
    try:
        with open(os.path.join(sys_pci, bus_id.lower(), "numa_node")) as f:
            node = int(f.read().strip())
    except (OSError, ValueError):
        return None
    return node if node >= 0 else None


def cpus_of_node(node, sys_node=_SYS_NODE):
    try:
        with open(os.path.join(sys_node, f"node{node}", "cpulist")) as f:
            return parse_cpulist(f.read())
    except (OSError, ValueError):
        return set()


def bind_to_device_node(device, sys_pci=_SYS_PCI, sys_node=_SYS_NODE):
    """Restrict the calling process to the CPUs of the NUMA node `device` hangs off.  Call it before allocating pinned memory
    and before starting threads.  Returns {"bound": bool, "node": int | None, "cpus": n, "reason": str}."""
    try:
        bus = device_pci_bus_id(device)
    except Exception as e:           # no device / library missing: nothing to bind to
        return {"bound": False, "node": None, "cpus": 0, "reason": f"no PCI id: {e}"}
    node = node_of_pci_device(bus, sys_pci)
    if node is None:
        return {"bound": False, "node": None, "cpus": 0, "reason": f"sysfs has no NUMA node for {bus}"}
    cpus = cpus_of_node(node, sys_node)
    allowed = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else set()
    target = cpus & allowed if allowed else cpus
    if not target:
        return {"bound": False, "node": node, "cpus": 0, "reason": f"node {node} has no CPU this process may use"}
    try:
        os.sched_setaffinity(0, target)
    except (OSError, AttributeError) as e:
        return {"bound": False, "node": node, "cpus": len(target), "reason": f"sched_setaffinity: {e}"}
    return {"bound": True, "node": node, "cpus": len(target), "reason": f"{bus} -> node {node}"}
