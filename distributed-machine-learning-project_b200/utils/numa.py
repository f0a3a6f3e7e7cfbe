"""Keep a rank's host threads (and therefore its pinned buffers) on the NUMA node of its GPU.

The end-to-end step is bound by the host -> device copy of its inputs; a pinned buffer
that lives on the other socket crosses the inter-socket link before it reaches PCIe.
The reference leaves placement to `mpirun --bind-to hwthread` (run_bench.sh:82-84); here
every rank binds itself before it allocates its staging buffers.  Everything in this module
is best effort: containers often hide the topology, and then nothing changes.
"""
import os


def _read(path):
    try:
        with open(path) as f:
            return f.read().strip()
    except OSError:
        return None


def parse_cpulist(text):
    """'0-3,8,10-11' -> {0, 1, 2, 3, 8, 10, 11}"""
    cpus = set()
    for part in (text or "").split(","):
        part = part.strip()
        if not part:
            continue
        lo, _, hi = part.partition("-")
        try:
            a, b = int(lo), int(hi) if hi else int(lo)
        except ValueError:
            return set()
        cpus.update(range(a, b + 1))
    return cpus


def gpu_numa_node(pci_bus_id, sys_root="/sys"):
    """NUMA node of a GPU given its PCI bus id ('0000:1B:00.0' or '00000000:1B:00.0'); None if unknown."""
    if not pci_bus_id:
        return None
    bus = pci_bus_id.lower()
    if len(bus.split(":")[0]) == 8:   # NVML prints an 8-digit domain, sysfs a 4-digit one
        bus = bus[4:]
    node = _read(os.path.join(sys_root, "bus/pci/devices", bus, "numa_node"))
    try:
        node = int(node)
    except (TypeError, ValueError):
        return None
    return node if node >= 0 else None


def node_cpus(node, sys_root="/sys"):
    return parse_cpulist(_read(os.path.join(sys_root, f"devices/system/node/node{node}/cpulist")))


def bind_to_gpu_node(device_index=0, sys_root="/sys"):
    """Restrict this process to the CPUs of the GPU's NUMA node (intersected with the CPUs it
    is allowed to use).  Returns the node number, or None when nothing was changed."""
    # opt in with DMLP_NUMA_BIND=1 until its effect on the copy rate has been measured on the target box
    if os.environ.get("DMLP_NUMA_BIND", "0") != "1" or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        import torch
        bus = torch.cuda.get_device_properties(device_index).pci_bus_id
        domain = getattr(torch.cuda.get_device_properties(device_index), "pci_domain_id", 0)
        dev_id = torch.cuda.get_device_properties(device_index).pci_device_id
        pci = f"{domain:04x}:{bus:02x}:{dev_id:02x}.0"
    except Exception:
        return None
    node = gpu_numa_node(pci, sys_root)
    if node is None:
        return None
    try:
        allowed = os.sched_getaffinity(0)
        want = node_cpus(node, sys_root) & allowed
        if len(want) < 2 or want == allowed:
            return None
        os.sched_setaffinity(0, want)
    except OSError:
        return None
    return node
