"""The multi-process island layout on CPU: world_size 2 over the gloo backend.

What is under test is the host-side logic the NCCL path uses on the GPUs -- the
global island numbering (island = local * world + rank), the exchange plan computed
by the C library (pgax_plan_exchange, a pure host function) and the order in which
sends and receives are posted.  The populations themselves are held by the CPU oracle
(this is a test: the product never computes on the CPU).  Each rank runs its islands
through tests/islands_host_schedule.run_islands_host with torch.distributed point-to-point
calls; the result must equal ONE oracle process running all islands.
"""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORLD = 2
SIZE, LENGTH, GENS, EVERY, PCT = 600, 256, 9, 2, 4.0


class OracleEngine:
    """Adapter: the engine interface of run_islands_host over oracle populations."""

    def __init__(self, po, oracle, pops):
        self.po, self.o, self.islands = po, oracle, pops

    def generation(self, pop):
        self.o.fill_random_values(pop)
        self.o.evaluate(pop)
        self.o.crossover(pop)
        self.o.mutate(pop)
        self.o.swap_generations(pop)
        self.o.evaluate(pop)

    def migration_count(self, pop, pct):
        return int(self.po.lib().opga_migration_count(pop.size, pct))

    def take_emigrants(self, pop, k):
        idx = self.po.select_top(pop.scores(), k)
        return np.concatenate([pop.genes()[idx].view(np.uint32).ravel(), pop.scores()[idx].view(np.uint32)])

    def put_immigrants(self, pop, k, payload):
        idx = self.po.select_top(pop.scores(), k, worst=True)
        rows, scores = pop.genes(), pop.scores()
        rows[idx] = payload[:k * pop.row_words].reshape(k, pop.row_words).view(rows.dtype)
        scores[idx] = payload[k * pop.row_words:].view(np.float32)
        pop.set_genes(rows)
        pop.set_scores(scores)


def worker(rank, port, k_local, topology, delay, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    from oracle import pyoracle as po
    from libpga_b200 import islands
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import islands_host_schedule
    o = po.Oracle(99)
    o.set_builtin_objective(po.OBJ_ONEMAX)
    pops = [o.create_population(SIZE, LENGTH, po.GENES_BITS, island=islands.island_id(s, rank, WORLD)) for s in range(k_local)]
    for p in pops:
        o.evaluate(p)
    eng = OracleEngine(po, o, pops)
    words = pops[0].row_words

    def send(payload, dst):
        dist.send(torch.from_numpy(payload.view(np.int32).copy()), dst)

    def recv(src):
        k = eng.migration_count(pops[0], PCT)
        t = torch.empty(k * (words + 1), dtype=torch.int32)
        dist.recv(t, src)
        return t.numpy().view(np.uint32)

    islands_host_schedule.run_islands_host(eng, GENS, EVERY, PCT, seed=99, topology=topology, delay=delay, rank=rank, world=WORLD,
                             send=send, recv=recv)
    for s, p in enumerate(pops):
        out[islands.island_id(s, rank, WORLD)] = (po.digest(p.genes()), po.digest(p.scores()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("k_local,topology,delay", [(1, 0, 1), (2, 0, 1), (2, 1, 0), (3, 1, 1)])
def test_two_ranks_equal_one_process(oracle, k_local, topology, delay):
    mgr = mp.Manager()
    out = mgr.dict()
    port = 29500 + (os.getpid() % 2000) + k_local * 7 + topology * 3 + delay
    mp.spawn(worker, args=(port, k_local, topology, delay, out), nprocs=WORLD, join=True)

    o = oracle.Oracle(99)
    o.set_builtin_objective(oracle.OBJ_ONEMAX)
    pops = [o.create_population(SIZE, LENGTH, oracle.GENES_BITS) for _ in range(k_local * WORLD)]
    o.set_topology(topology)
    o.set_migration_delay(delay)
    o.run_islands(GENS, EVERY, PCT)
    for g, p in enumerate(pops):
        assert out[g] == (oracle.digest(p.genes()), oracle.digest(p.scores())), f"island {g}"
    o.close()


def test_rendezvous_single_process():
    # world size 1: no communicator, the layout is the identity
    from libpga_b200 import islands
    assert islands.island_id(3, 0, 1) == 3
    assert islands.island_id(1, 1, 2) == 3
    assert islands.env_rank_world()[1] >= 1
