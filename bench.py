#!/usr/bin/env python
"""Benchmark of the generation hot path (pga_run / pga_run_islands).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload onemax|rastrigin|test1|tsp] [--gens G]

A "step" is one pga_run(G) call (G generations, default 100: the generation
count of the reference's own test/test.cu:43) on the workload BASELINE.json's
metric is quoted on: OneMax, 2^20 genomes x 1024 packed bits, one island per GPU.
With N > 1 (torchrun, one rank per GPU) every rank holds one such island, the
islands form a ring and the top 1 % migrate every 10 generations over NCCL
(pga_run_islands): weak scaling.

Prints ONE JSON line (rank 0).  value = genome-evaluations/s with the population
resident in HBM; e2e = the same through the public API with the initial
population uploaded from pinned host memory and the best individual read back
every step.  Two sub-records ride on the same line:
  c5               BASELINE config 5 as written: 8 islands x 2^22 x 1024 bits spread 8/N per
                   GPU, ring, top 1 % (41 943 rows) every 10 generations, 20 generations per
                   step, CUDA-event timed (skip with --no-c5)
  islands_parity   at N > 1: a small ring (one 20 000 x 1024-bit island per GPU, 23
                   generations, top 1 % every 5) run over NCCL and replayed on the CPU
                   oracle outside the timed region; ok = every island bit-identical
--impl reference times the CPU oracle (the reference has no CPU implementation of its own
and its GPU kernels cannot launch this workload: genome_len > 192, src/pga.cu:66-70) on
all host cores; at --gpus N it runs the same N-island ring (o.run_islands).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (size, genome_len, encoding, builtin objective / hooks, genome bytes)
    "onemax": dict(size=1 << 20, length=1024, bits=True, builtin="ONEMAX", label="C2 OneMax 2^20 x 1024 bits"),
    "rastrigin": dict(size=1 << 20, length=256, bits=False, builtin="RASTRIGIN", label="C3 Rastrigin 2^20 x 256 f32"),
    "test1": dict(size=40000, length=100, bits=False, hook="sum_obj", label="C1 test/test.cu 40000 x 100 f32, sum hook"),
    "tsp": dict(size=1 << 22, length=100, bits=False, hook="tsp_obj", xover="tsp_crossover",
                label="C4 test3 hooks 2^22 x 100 f32"),
}


KERNELS = {
    "onemax": "generation_fused_tiled<ENC_BITS, OBJ_ONEMAX, LOGG=3, NB=1, TB=8, EXACT, MINB=2, ONEBLOCK=0>",
    "rastrigin": "generation_fused_tiled<ENC_F32, OBJ_RASTRIGIN, LOGG=5, NB=2, TB=2, EXACT, MINB=3, ONEBLOCK=0>",
    "test1": "generation_hooks<PH_FUSED> (user objective through a device function pointer)",
    "tsp": "generation_hooks<CROSSOVER+MUTATE> then generation_hooks<EVALUATE> (user hooks through device function pointers; split chosen by timing)",
}


BOUND_NOTE = {
    "onemax": "516 B cross L2 per child (4 score-gather sectors + 2 parents + child) for 256 B algorithmic; latency-bound at "
              "4 warps per scheduler (3.8 of 9.5 warp-cycles per issue on the long scoreboard), no unit above 57 % of its peak; "
              "L2 hit rate 33 % since the row loads name the normal L2 policy, DRAM 169 MB read + 104 MB written per launch "
              "(ncu, profiles/r2_ncu_onemax.txt; profiles/r2_sweeps.md sections 1-2b)",
    "rastrigin": "3 population passes of DRAM traffic (2 gathered parents + 1 child) = 3.19 GB per launch at 93 % of the measured "
                 "peak, and 335 M warp instructions at 59 % issue efficiency: co-bound (profiles/r2_sweeps.md section 3)",
}


def genome_bytes(w):
    return w["length"] // 8 if w["bits"] else w["length"] * 4


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(workload):
    """Per-launch DRAM bytes of the fused kernel from the committed ncu capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f).get(workload, {}).get("dram_bytes_per_launch")
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks and throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []
        self.t_begin = self.t_end = None

    def mark_begin(self):
        self.t_begin = time.time()

    def mark_end(self):
        self.t_end = time.time()

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "25"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        # samples taken inside the timed region (with 0.15 s slack); if the region was
        # shorter than one sampling period fall back to everything since the warm-up
        inside = [l for (t, l) in self.lines if self.t_begin and self.t_end and self.t_begin - 0.15 <= t <= self.t_end + 0.15]
        for line in (inside or [l for (_, l) in self.lines]):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def configure(pg, g, w):
    if w.get("builtin"):
        g.set_builtin_objective(getattr(pg, "OBJ_" + w["builtin"]))
    if w.get("hook"):
        g.set_objective_function(w["hook"])
    if w.get("xover"):
        g.set_crossover_function(w["xover"])
    if w.get("hook") == "tsp_obj":
        import numpy as np
        with open(os.path.join(ROOT, "tests", "golden", "tsp100.txt")) as f:
            tok = f.read().split()
        n = int(tok[0])
        g.tsp_load(np.array(tok[1:1 + n * n], dtype=np.float32).reshape(n, n))


def run_reference_arm(args, w, emit):
    """The CPU oracle on all host cores: a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pyoracle as po
    cores = os.cpu_count() or 1
    enc = po.GENES_BITS if w["bits"] else po.GENES_F32
    # the same island layout as our arm: one island per GPU, ring, top pct every m generations
    n_islands = max(1, args.gpus) * args.islands
    ring = n_islands > 1
    o = po.Oracle(42, threads=cores)
    pops = [o.create_population(w["size"], w["length"], enc) for _ in range(n_islands)]
    if w.get("builtin"):
        o.set_builtin_objective(getattr(po, "OBJ_" + w["builtin"]))
    if w.get("hook"):
        o.set_objective_hook(w["hook"])
    if w.get("xover"):
        o.set_crossover_hook(w["xover"])
    if w.get("hook") == "tsp_obj":
        import numpy as np
        with open(os.path.join(ROOT, "tests", "golden", "tsp100.txt")) as f:
            tok = f.read().split()
        n = int(tok[0])
        o.tsp_load(np.array(tok[1:1 + n * n], dtype=np.float32).reshape(n, n))
    o.set_topology(po.TOPO_RING)
    o.evaluate_all()

    def step(gens):
        if ring:
            o.run_islands(gens, args.migrate_every, args.migrate_pct)
        else:
            o.run(gens)

    # size the per-step sample so the whole run stays within a few minutes
    t0 = time.perf_counter(); step(1); per_gen = time.perf_counter() - t0
    budget = 120.0 / max(1, args.steps + args.warmup)
    gens = max(1, min(args.gens, int(budget / max(per_gen, 1e-6))))
    if ring and gens > args.migrate_every:  # whole migration periods, so every step holds the same number of exchanges
        gens -= gens % args.migrate_every
    for _ in range(args.warmup):
        step(gens)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(gens)
    dt = time.perf_counter() - t0
    value = args.steps * gens * w["size"] * n_islands / dt
    sample = (f"{gens} generations of {w['label']} per step at full population size, {n_islands} island(s)"
              + (f", ring, top {args.migrate_pct}% every {args.migrate_every} generations" if ring else "") + ", OpenMP over individuals")
    o.close()
    line = {
        "impl": "reference", "metric": "genome-evaluations/sec", "value": value, "unit": "evals/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32" if w["bits"] else "f32", "data": "synthetic",
        "generations_per_sec": args.steps * gens / dt,
        "config": {"workload": w["label"], "generations_per_step": gens, "islands": n_islands, "islands_per_gpu": args.islands, "seed": 42,
                   "operators": "tournament-2, uniform crossover, p=0.01 single-gene mutation",
                   "migration": f"ring, top {args.migrate_pct}% every {args.migrate_every} generations (CPU)" if ring else "none (single island)",
                   "l2": "n/a (host memory)",
                   "note": "CPU oracle (oracle/pga_oracle.c): the reference has no CPU path and its own GPU kernels cannot launch this "
                           "workload (genome_len > 192, src/pga.cu:66-70); generations_per_step is a bounded sample of our arm's"},
        "cpu_baseline": {"value": value, "unit": "evals/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if n_islands == 1:
        rg = reference_gpu(w, args)  # test1 only: the reference's own GPU library runs this workload
        if rg:
            line["reference_gpu"] = rg
    emit(line)


def reference_gpu(w, args):
    """The reference's own GPU library (src/pga.cu rebuilt for sm_100a, oracle/build_ref.sh) and this
    repository's libpga.a behind the SAME program (oracle/ref_time.cu + the reference demo's objective),
    wall-clock pga_run on this box's GPU.  Only the reference's test/ workload: its kernels need
    64 * genome_len * 4 bytes of shared memory per block and have no packed-bit genomes."""
    import subprocess
    if w.get("hook") != "sum_obj" or w["length"] != 100:
        return None
    out = {}
    for name in ("ref1", "ours1"):
        exe = os.path.join(ROOT, "oracle", "_ref", "time", name)
        if not os.path.exists(exe):
            return {"unavailable": "oracle/_ref/time not built (oracle/build_ref.sh needs the reference checkout)"}
        r = subprocess.run([exe, "--size", str(w["size"]), "--gens", str(args.gens), "--reps", "5"],
                           capture_output=True, text=True, timeout=600)
        if r.returncode != 0:
            return {"unavailable": f"{name} exited {r.returncode}: {r.stderr[-200:]}"}
        out[name] = json.loads(r.stdout.strip().splitlines()[-1])
    ref, ours = out["ref1"], out["ours1"]
    return {"what": "pga_run wall clock, /root/reference/test/test.cu's workload through oracle/ref_time.cu: the unmodified reference "
                    "library rebuilt for sm_100a (cuRAND, time seed) against this repository's libpga.a, same program, same GPU",
            "reference": {"value": ref["evals_per_s"], "unit": "evals/s", "us_per_generation": ref["us_per_generation"]},
            "this_library": {"value": ours["evals_per_s"], "unit": "evals/s", "us_per_generation": ours["us_per_generation"]},
            "size": ref["size"], "genome_len": ref["genome_len"], "generations": ref["generations"]}


def cpu_baseline(w, cores=None):
    from oracle import pyoracle as po
    cores = cores or os.cpu_count() or 1
    enc = po.GENES_BITS if w["bits"] else po.GENES_F32
    o = po.Oracle(42, threads=cores)
    pop = o.create_population(w["size"], w["length"], enc)
    if w.get("builtin"):
        o.set_builtin_objective(getattr(po, "OBJ_" + w["builtin"]))
    if w.get("hook"):
        o.set_objective_hook(w["hook"])
    if w.get("xover"):
        o.set_crossover_hook(w["xover"])
    if w.get("hook") == "tsp_obj":
        import numpy as np
        with open(os.path.join(ROOT, "tests", "golden", "tsp100.txt")) as f:
            tok = f.read().split()
        n = int(tok[0])
        o.tsp_load(np.array(tok[1:1 + n * n], dtype=np.float32).reshape(n, n))
    o.evaluate(pop)
    t0 = time.perf_counter(); o.run(1); per_gen = time.perf_counter() - t0
    gens = max(1, min(100, int((15.0 if cores > 1 else 6.0) / max(per_gen, 1e-6))))
    t0 = time.perf_counter(); o.run(gens); dt = time.perf_counter() - t0
    o.close()
    return {"value": gens * w["size"] / dt, "unit": "evals/s", "cores": cores, "kind": "port",
            "sample": f"{gens} generations at full population size ({w['label']}), oracle/pga_oracle.c with OpenMP over individuals, {dt:.1f} s"}


def run_c5(args, pg, islands, dist, torch, stream, world, distributed):
    """BASELINE config 5 as written: 8 islands x 2^22 x 1024 bits over the GPUs, ring, top 1 %
    (41 943 rows) every 10 generations; 20 generations per step; CUDA events, max over ranks."""
    per_gpu = max(1, 8 // world)
    size, bits, gens, every, pct = 1 << 22, 1024, 20, 10, 1.0
    g = pg.PGA(seed=42)
    g.set_stream(stream.cuda_stream)
    if distributed:
        islands.rendezvous(g, dist)
    g.set_topology(pg.TOPO_RING)
    pops = [g.create_population(size, bits, encoding=pg.GENES_BITS) for _ in range(per_gpu)]
    g.set_builtin_objective(pg.OBJ_ONEMAX)
    g.evaluate_all()
    for _ in range(3):
        g.run_islands(gens, every, pct)
    if distributed:
        dist.barrier()
    torch.cuda.synchronize()
    steps = 3
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = g.kernel_launches
    e0.record(stream)
    for _ in range(steps):
        g.run_islands(gens, every, pct)
    e1.record(stream)
    if distributed:
        dist.barrier()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if distributed:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    launches = g.kernel_launches - launches0
    assert pg.migration_count(size, pct) == 41943
    g.deinit()
    total_islands = per_gpu * world
    return {"workload": "BASELINE config 5: %d islands x 2^22 x 1024 bits, ring, top 1%% (41943 rows) every 10 generations" % total_islands,
            "value": steps * gens * size * total_islands / (ms * 1e-3), "unit": "evals/s", "ms_per_step": ms / steps,
            "generations_per_step": gens, "steps": steps, "islands": total_islands, "islands_per_gpu": per_gpu, "n_gpus": world,
            "exchanges_per_step": 1, "gpu_launches": launches}


def run_islands_parity(pg, islands, dist, rank, world):
    """One 20 000 x 1024-bit island per GPU, ring over NCCL, 23 generations, top 1 % every 5;
    rank 0 replays all islands on the CPU oracle and compares digests island by island."""
    from oracle import pyoracle as po
    seed, size, bits, gens, every, pct = 123, 20000, 1024, 23, 5, 1.0
    g = pg.PGA(seed=seed)
    islands.rendezvous(g, dist)
    g.set_topology(pg.TOPO_RING)
    pop = g.create_population(size, bits, encoding=pg.GENES_BITS)
    g.set_builtin_objective(pg.OBJ_ONEMAX)
    g.run_islands(gens, every, pct)
    mine = {pop.island: (po.digest(pop.genes()), po.digest(pop.scores()))}
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    g.deinit()
    if rank != 0:
        return None
    got = {}
    for d in gathered:
        got.update(d)
    o = po.Oracle(seed, threads=os.cpu_count() or 1)
    o.set_builtin_objective(po.OBJ_ONEMAX)
    ops = [o.create_population(size, bits, po.GENES_BITS) for _ in range(world)]
    o.set_topology(po.TOPO_RING)
    o.run_islands(gens, every, pct)
    bad = [i for i, p in enumerate(ops) if got.get(i) != (po.digest(p.genes()), po.digest(p.scores()))]
    o.close()
    return {"checked": True, "ok": not bad, "islands": world, "mismatched_islands": bad,
            "what": "one 20000 x 1024-bit island per GPU, NCCL ring, 23 generations, top 1% every 5; every island's genomes and scores "
                    "compared bit for bit with oracle/pga_oracle.c (outside the timed region)"}


def main():
    # the contract is ONE JSON line on stdout: libraries (NCCL prints its version banner
    # to stdout) must not get there, so fd 1 points at stderr until the line is ready
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        os.write(real_stdout, (json.dumps(line) + "\n").encode())

    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="onemax", choices=sorted(WORKLOADS))
    ap.add_argument("--gens", type=int, default=100, help="generations per step")
    ap.add_argument("--size", type=int, default=0, help="override the population size")
    ap.add_argument("--length", type=int, default=0, help="override the genome length (sweeps only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--islands", type=int, default=1, help="islands per GPU (BASELINE config 5: 8 islands over N GPUs)")
    ap.add_argument("--migrate-every", type=int, default=10)
    ap.add_argument("--migrate-pct", type=float, default=1.0)
    ap.add_argument("--no-c5", action="store_true", help="skip the BASELINE config 5 sub-record")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    w = dict(WORKLOADS[args.workload])
    if args.size:
        w["size"] = args.size
        w["label"] += f" (size overridden to {args.size})"
    if args.length:
        w["length"] = args.length
        w["label"] += f" (length overridden to {args.length})"

    if args.impl == "reference":
        run_reference_arm(args, w, emit)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    import libpga_b200 as pg
    from libpga_b200 import islands

    rank, world, local = islands.env_rank_world()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local)
    distributed = world > 1
    if distributed:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    enc = pg.GENES_BITS if w["bits"] else pg.GENES_F32
    stream = torch.cuda.current_stream()
    g = pg.PGA(seed=42)
    g.set_stream(stream.cuda_stream)
    if distributed:
        islands.rendezvous(g, dist)
        g.set_topology(pg.TOPO_RING)
    pops = [g.create_population(w["size"], w["length"], encoding=enc) for _ in range(args.islands)]
    pop = pops[0]
    configure(pg, g, w)
    if args.islands > 1 and not distributed:
        g.set_topology(pg.TOPO_RING)

    def step():
        if distributed or args.islands > 1:
            g.run_islands(args.gens, args.migrate_every, args.migrate_pct)
        else:
            g.run(args.gens)

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    sampler.start()
    g.evaluate_all()
    for _ in range(args.warmup):
        step()

    # ---- value: population resident in HBM, CUDA events on the launching stream
    barrier()
    launches0 = g.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms = 0.0
    barrier()
    sampler.mark_begin()
    e0.record(stream)
    for _ in range(args.steps):
        step()
        kernel_ms += g.last_run_ms
    e1.record(stream)
    barrier()
    sampler.mark_end()
    elapsed_ms = e0.elapsed_time(e1)
    launches = g.kernel_launches - launches0
    if distributed:
        t = torch.tensor([elapsed_ms, kernel_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms, kernel_ms = float(t[0]), float(t[1])
    evals = args.steps * args.gens * w["size"] * world * args.islands
    value = evals / (elapsed_ms * 1e-3)

    # ---- e2e: initial population from pinned host memory, best individual back, every step
    host_rows = torch.empty((w["size"], pop.row_words), dtype=torch.int32, pin_memory=True)
    host_np = host_rows.numpy().view(np.uint32 if w["bits"] else np.float32)
    pop.genes(out=host_np)
    h2d = host_np.nbytes
    d2h = w["length"] * 4 + 16
    e2e_steps = max(1, args.steps)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        pop.set_genes(host_np)          # H2D inside the call
        step()
        row = g.get_best(pop)           # D2H inside the call
    barrier()
    e2e_s = time.perf_counter() - t0
    if distributed:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t[0])
    # +1: the uploaded generation is scored too (only island 0 is re-uploaded per step)
    e2e_value = e2e_steps * (args.gens * args.islands + 1) * w["size"] * world / e2e_s
    clocks = sampler.stop()

    peak, peak_src = measured_peak()
    gb = genome_bytes(w)
    n_launch = args.steps * args.gens * args.islands
    achieved = (2.0 * w["size"] * gb) / (kernel_ms * 1e-3 / n_launch) / 1e9  # algorithmic GB/s of the fused kernel
    line = {
        "metric": "genome-evaluations/sec", "value": value, "unit": "evals/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32" if w["bits"] else "f32", "data": "synthetic",
        "generations_per_sec": args.steps * args.gens / (elapsed_ms * 1e-3),
        "config": {"workload": w["label"], "generations_per_step": args.gens, "islands": world * args.islands, "islands_per_gpu": args.islands, "seed": 42,
                   "operators": "tournament-2, uniform crossover, p=0.01 single-gene mutation",
                   "migration": f"ring, top {args.migrate_pct}% every {args.migrate_every} generations over NCCL" if (distributed or args.islands > 1) else "none (single island)",
                   "l2": "inputs larger than L2: two generations of the population = %.0f MiB vs 126 MB" % (2 * w["size"] * gb / 2 ** 20),
                   "note": "CUDA library through the C-ABI; no CPU path"},
        "e2e": {"value": e2e_value, "unit": "evals/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "what": "upload population (pinned host) + pga_run/pga_run_islands + pga_get_best, wall clock"},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(args.workload), "kernel": KERNELS[args.workload],
                     "algorithmic_bytes_per_launch": 2 * w["size"] * gb, "avg_launch_ms": kernel_ms / n_launch,
                     "peak_source": peak_src, "frac_of_8TBs_nominal": achieved / 8000.0,
                     "note": BOUND_NOTE.get(args.workload)},
    }
    g.deinit()
    if args.workload == "onemax" and not args.no_c5 and not args.size and not args.length:
        c5 = run_c5(args, pg, islands, dist, torch, stream, world, distributed)
        line["c5"] = c5
    if distributed:
        parity = run_islands_parity(pg, islands, dist, rank, world)
        line["islands_parity"] = parity
    else:
        line["islands_parity"] = {"checked": False, "why": "single GPU: no NCCL exchange to check (tests/test_gpu_islands.py covers the one-GPU islands)"}
    if rank == 0 and not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_baseline(w)                  # all host cores (OpenMP over individuals)
        rg = reference_gpu(w, args)                             # the reference's own GPU library, where it can run
        if rg:
            line["reference_gpu"] = rg
        line["cpu_baseline_serial"] = cpu_baseline(w, cores=1)  # the serial oracle, one core
    if distributed:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        emit(line)


if __name__ == "__main__":
    main()
