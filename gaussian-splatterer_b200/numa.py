"""Host-memory placement for the truth-frame uploads: a pinned staging block should live on the NUMA node the GPU's
PCIe root hangs off, otherwise every upload crosses the socket interconnect (and, with one process per GPU started from
the same shell, all blocks pile up on node 0).  Linux only; every function degrades to a no-op."""
import ctypes
import os

_SYS_set_mempolicy = 238          # x86_64
MPOL_DEFAULT, MPOL_PREFERRED, MPOL_BIND = 0, 1, 2


def gpu_numa_node(pci_bus_id):
    """NUMA node of a GPU from sysfs (pci_bus_id like '00000000:1B:00.0'); None when unknown."""
    if not pci_bus_id:
        return None
    bdf = pci_bus_id.lower()
    if len(bdf.split(":")[0]) == 8:
        bdf = bdf[4:]
    try:
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
        return node if node >= 0 else None
    except Exception:
        return None


def node_cpus(node):
    try:
        txt = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
    except Exception:
        return []
    cpus = []
    for part in txt.split(","):
        if "-" in part:
            a, b = part.split("-")
            cpus.extend(range(int(a), int(b) + 1))
        elif part:
            cpus.append(int(part))
    return cpus


def set_mempolicy(mode, node=None):
    """set_mempolicy(2) for the calling thread; returns True on success."""
    try:
        libc = ctypes.CDLL(None, use_errno=True)
        if node is None:
            return libc.syscall(_SYS_set_mempolicy, MPOL_DEFAULT, None, 0) == 0
        mask = ctypes.c_ulong(1 << node)
        return libc.syscall(_SYS_set_mempolicy, mode, ctypes.byref(mask), ctypes.sizeof(mask) * 8) == 0
    except Exception:
        return False


class local_to_gpu:
    """Context manager: host allocations made inside prefer the GPU's NUMA node, and the thread runs on that node's
    CPUs where the cpuset allows it.  `info` records what was done (for the bench JSON)."""

    def __init__(self, pci_bus_id):
        self.node = gpu_numa_node(pci_bus_id)
        self.info = {"pci_bus_id": pci_bus_id, "numa_node": self.node, "mempolicy": False, "affinity": False}
        self._old_aff = None

    def __enter__(self):
        if self.node is None:
            return self
        self.info["mempolicy"] = set_mempolicy(MPOL_PREFERRED, self.node)
        try:
            cpus = set(node_cpus(self.node)) & set(os.sched_getaffinity(0))
            if cpus:
                self._old_aff = os.sched_getaffinity(0)
                os.sched_setaffinity(0, cpus)
                self.info["affinity"] = True
        except Exception:
            pass
        return self

    def __exit__(self, *exc):
        if self.node is not None:
            set_mempolicy(MPOL_DEFAULT)
            if self._old_aff is not None:
                try:
                    os.sched_setaffinity(0, self._old_aff)
                except Exception:
                    pass
        return False
