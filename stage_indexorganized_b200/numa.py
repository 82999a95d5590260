"""Host placement for the host-buffer (e2e) path: pin the calling process to the CPUs of the NUMA node its GPU hangs
off, so that the pinned result buffers it allocates afterwards are local to that GPU's PCIe root (first-touch policy).
Best effort: returns None where the topology is not exposed (e.g. a VM without NUMA information)."""
import os


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_numa_node(device_index):
    import torch
    p = torch.cuda.get_device_properties(device_index)
    bdf = f"{getattr(p, 'pci_domain_id', 0):04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    try:
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            return int(f.read()), bdf
    except (OSError, ValueError):
        return -1, bdf


def bind_to_gpu_node(device_index):
    """-> {"node", "cpus", "bdf"} after sched_setaffinity to the GPU's node, or None if nothing was done."""
    node, bdf = gpu_numa_node(device_index)
    if node < 0:
        return None
    try:
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = _parse_cpulist(f.read())
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"node": node, "cpus": len(cpus), "bdf": bdf}
    except OSError:
        return None
