"""Process-per-GPU island layout: rendezvous and the exchange schedule.

The schedule itself is computed by the C library (pgax_plan_exchange, a pure host
function), so what is tested on CPU with gloo is the same code the NCCL path
follows on the GPUs.  This module adds the torch.distributed plumbing around it:

* rendezvous(): rank 0 creates the NCCL unique id inside the C library, the bytes
  travel through torch.distributed, every rank hands them to pgax_dist_init.
The pga_run_islands schedule itself runs natively in the C library (pga_host.c);
tests/islands_host_schedule.py restates it over torch.distributed point-to-point
calls so that the multi-rank bookkeeping can be exercised on CPU with gloo.
"""
import ctypes as C
import os

from . import _lib

MAX_GLOBAL_ISLANDS = 640


class Transfer(C.Structure):
    _fields_ = [("src_island", C.c_uint32), ("dst_island", C.c_uint32),
                ("src_rank", C.c_int32), ("dst_rank", C.c_int32),
                ("src_local", C.c_uint32), ("dst_local", C.c_uint32)]


def island_id(local_index, rank, world):
    """Global island number of local population `local_index` on `rank`."""
    return local_index * world + rank


def plan_exchange(seed, round_, topology, world, islands_per_rank):
    """Transfers of exchange number `round_`, ascending source-island order."""
    L = _lib.load()
    buf = (Transfer * MAX_GLOBAL_ISLANDS)()
    n = L.pgax_plan_exchange(seed, round_, topology, world, islands_per_rank, buf, MAX_GLOBAL_ISLANDS)
    if n < 0:
        raise ValueError(f"pgax_plan_exchange failed: {n}")
    return [dict(src_island=t.src_island, dst_island=t.dst_island, src_rank=t.src_rank, dst_rank=t.dst_rank,
                 src_local=t.src_local, dst_local=t.dst_local) for t in buf[:n]]


def env_rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def rendezvous(pga, dist=None):
    """Give every rank's solver the same NCCL communicator.  `dist` is an
    initialised torch.distributed module (any backend); with world size 1 nothing
    is exchanged."""
    from . import api
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        pga.dist_set_layout(0, 1)
        return 0, 1
    rank, world = dist.get_rank(), dist.get_world_size()
    box = [api.dist_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    pga.dist_init(rank, world, box[0])
    return rank, world
