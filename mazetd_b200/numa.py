"""Host-side placement helper: run the calling process on the CPUs of the NUMA node its GPU hangs off, so that pinned host
buffers allocated afterwards (first touch / cudaHostAlloc follow the allocating thread) sit on that node and device->host
copies do not cross the socket interconnect. Used by bench.py's e2e leg and tools/d2h_bench.py at N > 1; a no-op when the
topology cannot be read (containers without /sys, single-node hosts)."""
import os


def _read(path):
    try:
        with open(path) as f:
            return f.read().strip()
    except OSError:
        return None


def _parse_cpulist(s):
    cpus = set()
    for part in filter(None, (s or "").split(",")):
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def gpu_pci_bus_id(device_index):
    """'0000:1b:00.0' style id of a CUDA device (lower case, as /sys spells it), or None"""
    try:
        import torch

        p = torch.cuda.get_device_properties(device_index)
        dom = getattr(p, "pci_domain_id", 0)
        return "%04x:%02x:%02x.0" % (dom, p.pci_bus_id, p.pci_device_id)
    except Exception:
        pass
    try:
        import pynvml

        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[device_index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else device_index
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(idx)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        return bus.lower()[-12:]
    except Exception:
        return None


def bind_to_gpu_node(device_index):
    """returns {'node': n, 'cpus': count, 'bound': bool}; leaves the affinity alone when anything is unknown"""
    info = {"node": None, "cpus": len(os.sched_getaffinity(0)), "bound": False}
    if os.environ.get("MTD_NO_NUMA_BIND"):
        return info
    bus = gpu_pci_bus_id(device_index)
    if not bus:
        return info
    node = _read("/sys/bus/pci/devices/%s/numa_node" % bus)
    if node is None or int(node) < 0:
        return info
    info["node"] = int(node)
    cpus = _parse_cpulist(_read("/sys/devices/system/node/node%d/cpulist" % int(node)))
    allowed = os.sched_getaffinity(0)
    cpus &= allowed
    if not cpus or cpus == allowed:
        return info
    try:
        os.sched_setaffinity(0, cpus)
    except OSError:
        return info
    info["cpus"] = len(cpus)
    info["bound"] = True
    return info
