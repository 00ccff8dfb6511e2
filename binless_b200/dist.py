"""One process per GPU: joins the library's NCCL communicator from a ``torch.distributed`` process group.

After ``init_comm()`` every fused-lasso solve made through the C ABI is shared by the ranks (trail-block sharding,
include/fastbinless.h: fb_comm_init); every rank must then make the same calls with the same inputs. torch is only
the plumbing: it hands the 128-byte NCCL id from rank 0 to the others (any backend: nccl on a GPU box, gloo on CPU).
"""
from __future__ import annotations

import os


def _nccl_library_path():
    """The libnccl torch itself uses (nvidia-nccl wheel), so that the library and torch agree on one NCCL."""
    try:
        import nvidia.nccl as m
        base = list(m.__path__)[0]
        p = os.path.join(base, "lib", "libnccl.so.2")
        return p if os.path.exists(p) else None
    except Exception:
        return None


def init_comm(lib=None):
    """Collective over the default process group. Returns (rank, world)."""
    import torch
    import torch.distributed as dist
    from . import cuda_lib
    lib = lib or cuda_lib()
    if not dist.is_initialized() or dist.get_world_size() == 1:
        lib.comm_init(0, 1)
        return 0, 1
    rank, world = dist.get_rank(), dist.get_world_size()
    if "FB_NCCL_LIB" not in os.environ:
        p = _nccl_library_path()
        if p:
            os.environ["FB_NCCL_LIB"] = p
    uid = [lib.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    lib.comm_init(rank, world, uid[0])
    return rank, world


def shutdown_comm(lib=None):
    from . import cuda_lib
    (lib or cuda_lib()).comm_destroy()


def bind_to_gpu_numa_node(device_index: int):
    """Pin this process (and the library's staging threads, which inherit the mask) to the CPUs of the NUMA node the GPU
    hangs off, so that pinned buffers are first-touched there and the PCIe traffic of eight ranks does not cross sockets.
    Best effort: returns the node, or None when the topology cannot be read."""
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if allowed:
            os.sched_setaffinity(0, allowed)
            return node
    except Exception:
        pass
    return None
