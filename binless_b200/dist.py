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
