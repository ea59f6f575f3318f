"""Run under torchrun on N >= 2 GPUs (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/nccl_islands_worker.py

Every rank drives its islands through pga_run_islands (NCCL send/recv ring or random
topology inside the C library); rank 0 then replays ALL islands on the CPU oracle and
compares digests of genomes and scores island by island.  Prints NCCL-ISLANDS-OK."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import libpga_b200 as pg  # noqa: E402
from libpga_b200 import islands  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

SEED, SIZE, BITS, GENS, EVERY, PCT = 123, 20000, 1024, 23, 5, 1.0


def main():
    rank, world, local = islands.env_rank_world()
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    failures = 0
    # the last case arms the early stop (include/pga.h:147): every rank must leave the loop after
    # the same generation, which takes an all-reduce of the best score over the ranks
    for k_local, topology, delay, target in ((1, pg.TOPO_RING, 1, None), (2, pg.TOPO_RING, 0, None), (2, pg.TOPO_RANDOM, 1, None),
                                             (1, pg.TOPO_RING, 1, 590.0)):
        g = pg.PGA(seed=SEED)
        islands.rendezvous(g, dist)
        g.set_topology(topology)
        g.set_migration_delay(delay)
        pops = [g.create_population(SIZE, BITS, encoding=pg.GENES_BITS) for _ in range(k_local)]
        g.set_builtin_objective(pg.OBJ_ONEMAX)
        assert [p.island for p in pops] == [islands.island_id(s, rank, world) for s in range(k_local)]
        gens = GENS if target is None else 200
        if target is not None:
            g.set_target(target, 3)
        g.run_islands(gens, EVERY, PCT)
        ran = g.generations_run
        mine = {p.island: (po.digest(p.genes()), po.digest(p.scores())) for p in pops}
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        g.deinit()
        if rank == 0:
            got = {}
            for d in gathered:
                got.update(d)
            o = po.Oracle(SEED, threads=os.cpu_count() or 1)
            o.set_builtin_objective(po.OBJ_ONEMAX)
            ops = [o.create_population(SIZE, BITS, po.GENES_BITS) for _ in range(k_local * world)]
            o.set_topology(topology)
            o.set_migration_delay(delay)
            if target is not None:
                o.set_target(target, 3)
            o.run_islands(gens, EVERY, PCT)
            if target is not None:
                ok = ran == o.generations_run and 0 < ran < gens
                failures += 0 if ok else 1
                print(f"early stop: ran {ran} generations, oracle {o.generations_run}: {'ok' if ok else 'MISMATCH'}", flush=True)
            for i, p in enumerate(ops):
                ok = got[i] == (po.digest(p.genes()), po.digest(p.scores()))
                failures += 0 if ok else 1
                print(f"islands/rank={k_local} topology={topology} delay={delay} island {i}: {'ok' if ok else 'MISMATCH'}", flush=True)
            o.close()
        dist.barrier()
    if rank == 0:
        print("NCCL-ISLANDS-OK" if failures == 0 else f"NCCL-ISLANDS-FAILED ({failures})", flush=True)
    dist.destroy_process_group()
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
