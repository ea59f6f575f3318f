"""Test scaffolding: the pga_run_islands schedule (reference stub: src/pga.cu:393-395;
semantics: include/pga.h:145-150) written against an abstract engine and
torch.distributed point-to-point calls.  The product does NOT use it -- pga_run_islands
runs the schedule natively over NCCL (libpga_b200/csrc/pga_host.c) -- it exists so the
multi-rank bookkeeping (global island numbering, the exchange plan computed by the C
library's pgax_plan_exchange, the posting order of sends and receives) can be exercised
with the gloo backend on CPU."""
from libpga_b200.islands import plan_exchange


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
