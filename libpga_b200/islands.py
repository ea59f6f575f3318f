"""Process-per-GPU island layout: rendezvous and the exchange schedule.

The schedule itself is computed by the C library (pgax_plan_exchange, a pure host
function), so what is tested on CPU with gloo is the same code the NCCL path
follows on the GPUs.  This module adds the torch.distributed plumbing around it:

* rendezvous(): rank 0 creates the NCCL unique id inside the C library, the bytes
  travel through torch.distributed, every rank hands them to pgax_dist_init.
* run_islands_host(): the pga_run_islands schedule (reference stub:
  src/pga.cu:393-395; semantics: include/pga.h:145-150) written against an
  abstract engine and torch.distributed point-to-point calls.  The GPU product
  does NOT use it (pga_run_islands runs the schedule natively over NCCL); it
  exists so the multi-rank bookkeeping can be exercised with the gloo backend.
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


def run_islands_host(engine, n, m, pct, *, seed, topology, delay, rank, world, send, recv):
    """pga_run_islands over an abstract engine.

    engine: .islands -> list of local populations; .generation(pop);
            .migration_count(pop, pct); .take_emigrants(pop, k) -> payload;
            .put_immigrants(pop, k, payload)
    send(payload, dst_rank) / recv(src_rank) -> payload move one island's emigrants.
    """
    K = len(engine.islands)
    total = K * world
    pending = None
    round_ = 0
    for g in range(1, n + 1):
        for pop in engine.islands:
            engine.generation(pop)
        if pending is not None and pending["land"] == g:
            _finish(engine, pending)
            pending = None
        if m and g % m == 0 and g + delay <= n and total > 1 and pending is None:
            plan = plan_exchange(seed, round_, topology, world, K)
            round_ += 1
            pending = _start(engine, plan, pct, rank, send, recv)
            pending["land"] = g + delay
            if delay == 0:
                _finish(engine, pending)
                pending = None


def _start(engine, plan, pct, rank, send, recv):
    out = {}
    for t in plan:  # snapshot every local source first
        if t["src_rank"] == rank:
            src = engine.islands[t["src_local"]]
            k = engine.migration_count(src, pct)
            out[t["src_local"]] = (k, engine.take_emigrants(src, k) if k else None)
    arrivals = {}
    for t in plan:  # same order on every rank
        s_here, d_here = t["src_rank"] == rank, t["dst_rank"] == rank
        if s_here and d_here:
            arrivals[t["dst_local"]] = out[t["src_local"]]
        elif s_here:
            k, payload = out[t["src_local"]]
            if k:
                send(payload, t["dst_rank"])
        elif d_here:
            dst = engine.islands[t["dst_local"]]
            k = engine.migration_count(dst, pct)
            arrivals[t["dst_local"]] = (k, recv(t["src_rank"]) if k else None)
    return {"arrivals": arrivals}


def _finish(engine, pending):
    for local, (k, payload) in sorted(pending["arrivals"].items()):
        if k:
            engine.put_immigrants(engine.islands[local], k, payload)
