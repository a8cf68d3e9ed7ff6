"""Host-side placement for the host-buffer entry points (wg_commit_trace, wg_sharded_commit, wg_sharded_download).

Those calls are PCIe-bound (33 GB of LDE + tree per 2^22 x 100 commit come back to the host), and on a two-socket host a
pinned buffer that lives on the other socket's memory is reached through the inter-socket link: every rank of a multi-GPU
job then shares that one link. Pinned pages are placed on the NUMA node of the thread that allocates them, so the caller
binds itself to the GPU's node BEFORE allocating its pinned buffers. The reference has no counterpart (it has no device);
a Rust prover would do the same with `libnuma` or `sched_setaffinity` before `cudaHostAlloc`.

Plain sysfs reads + os.sched_setaffinity; no-ops (returning None) where the platform gives no answer."""
import os


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def gpu_numa_node(device_index):
    """NUMA node of CUDA device `device_index` (sysfs `numa_node` of its PCI function), or None."""
    try:
        import torch

        p = torch.cuda.get_device_properties(device_index)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            node = int(f.read().strip())
        return node if node >= 0 else None
    except Exception:
        return None


def node_cpus(node):
    try:
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            return _parse_cpulist(f.read())
    except Exception:
        return set()


class bind_to_gpu_node:
    """Context manager: restrict the calling process to the CPUs of the GPU's NUMA node (intersected with the CPUs it may
    already use) while pinned buffers are allocated and the host-buffer calls run; restores the previous affinity on exit.
    `.info` says what was done."""

    def __init__(self, device_index):
        self.device_index = device_index
        self.prev = None
        self.info = {"bound": False}

    def __enter__(self):
        try:
            prev = os.sched_getaffinity(0)
        except Exception:
            return self
        node = gpu_numa_node(self.device_index)
        if node is None:
            self.info["why"] = "no NUMA node reported for the device"
            return self
        cpus = node_cpus(node) & prev
        if not cpus or cpus == prev:
            self.info.update({"node": node, "why": "single node or no CPU of the node is usable"})
            return self
        os.sched_setaffinity(0, cpus)
        self.prev = prev
        self.info = {"bound": True, "node": node, "cpus": len(cpus)}
        return self

    def __exit__(self, *exc):
        if self.prev is not None:
            os.sched_setaffinity(0, self.prev)
        return False
