#!/usr/bin/env python
"""bench.py — Mutations-action queries/s on synthetic SARS-CoV-2-shaped data (BASELINE.json, configuration C5).

Workload: ONE database of --rows (default 40 M) synthetic sequences x 29,903 nt in lineage DatabasePartitions
(seed 0x51100005), sharded over the N ranks by DatabasePartition (strong scaling: N = 1 holds all of it, 4 GB of index).
For N > 1 the database is the same rows and values; by default every lineage partition is cut into N DatabasePartitions of
consecutive rows, one per GPU, so that every GPU holds 1/N of every lineage whatever the filter selects (--shard partitions
keeps the 23 whole partitions and assigns them by size: a filter that selects 7 of them leaves most GPUs idle).
One "step" = one nucleotide Mutations query (minProportion 0.05) out of a rotation of 8 distinct N-Of(3 of 10) filters over
clade-defining NucleotideEquals (different picks -> different sets of partial partitions, different |F|); step i runs
query i mod 8. For N > 1 every rank answers the query on its partitions and the count tables are summed inside the query's
submission (ncclReduce over NVLink, valid-symbol rows only); rank 0 selects the rows.

    value  queries/s with the queries prepared (plans resident in HBM, one CUDA graph launch per query at N = 1), CUDA
           events around the timed steps, max over ranks
    e2e    queries/s through the call a user makes: query JSON in -> response JSON out, plan cache OFF, i.e. every step
           parses, compiles FilterExpression -> Operator per partition, lowers, uploads the plans (H2D), runs, reads the
           rows back (D2H) and formats; e2e.cached is the same with the engine's plan cache on
    --impl reference   the CPU restatement of the reference's algorithm (oracle/, all host threads): the reference itself
           cannot be built here. Each step runs query 0 on a fixed sample of partitions whose share of the
           whole query's CPU time was measured by one full pass during set-up; ms_per_step is the time that was measured.
    --weak replicas of one --rows shard per rank (round-1 mode), --lineage adds the C5 lineage / True queries to the line.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
import zlib
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT / "lapis-silo_b200"))

GENOME_LENGTH = 29903
SIGMA = 16
VALID_ROWS = 5            # symbols -, A, C, G, T: the rows of the table anyone reads (and the rows that are reduced)
METRIC = "mutations_queries_per_s"
UNIT = "queries/s"
N_ROTATION = 8


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=int(os.environ.get("SILO_BENCH_ROWS", 40_000_000)))
    ap.add_argument("--seed", type=lambda s: int(s, 0), default=0x51100005)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--full-pass", action="store_true", help="--impl reference: measure the sample's share of the query by one full pass of the CPU restatement (minutes at 40 M rows)")
    ap.add_argument("--weak", action="store_true", help="N > 1: every rank holds its own shard of --rows sequences (replicated shards) instead of a share of one database")
    ap.add_argument("--shard", default="rows", choices=["rows", "partitions"],
                    help="N > 1, one database: 'rows' cuts every lineage partition into N DatabasePartitions of consecutive rows, one per GPU (every GPU holds 1/N of every lineage); "
                         "'partitions' assigns the 23 whole lineage partitions to the GPUs (longest-processing-time-first on rows)")
    ap.add_argument("--lineage", action="store_true", help="also measure the C5 lineage and True queries (full / partial / empty partitions)")
    ap.add_argument("--in-flight", type=int, default=8, help="queries in flight of the extra throughput measurement (N = 1 only; 0 = skip)")
    return ap.parse_args()


def nof_query(clade, k):
    """rotation member k: 3 of 10 clade-defining mutations, picks (7 j + 3 k) mod |clade|; k = 0 is the C3 query of round 1"""
    picks = [clade[(7 * j + 3 * k) % len(clade)] for j in range(10)]
    return {"action": {"type": "Mutations", "minProportion": 0.05},
            "filterExpression": {"type": "N-Of", "numberOfMatchers": 3, "matchExactly": False,
                                 "children": [{"type": "NucleotideEquals", "position": p + 1, "symbol": s} for p, s in picks]}}


def c3_query(db):
    return nof_query(db.clade_mutations(), 0)


def rotation(db):
    clade = db.clade_mutations()
    return [nof_query(clade, k) for k in range(N_ROTATION)]


def workload_name(rows):
    return f"C5: {rows} synthetic sequences x {GENOME_LENGTH} nt in lineage partitions, nucleotide Mutations under a rotation of {N_ROTATION} N-Of(3 of 10) filters"


class ClockSampler:
    """nvidia-smi clock / throttle-reason samples during the timed region (B200_PROFILING.md)."""
    FIELDS = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.samples, self.stop, self.thread = index, [], threading.Event(), None

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.splitlines()[0].split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()
        return self

    def __exit__(self, *exc):
        self.stop.set()
        self.thread.join(timeout=6)

    def summary(self):
        sm = [float(s[0]) for s in self.samples if s and s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if len(s) > 1 and s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for s in self.samples for i in range(4) if len(s) > 2 + i and s[2 + i].lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(self.samples)}


def measured_peak():
    path = ROOT / "MEASURED_PEAKS.json"
    if path.exists():
        return float(json.loads(path.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_calibrate(db, query, budget_s=3.0):
    """One full pass of the oracle (faithful mode = the reference's loops, all host threads) over every partition of this
    process, one partition at a time, timed per partition. Returns (full_seconds, sample partitions, sample_seconds): the
    sample is a prefix of a fixed pseudo-random order whose measured time reaches about budget_s."""
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_bind
    cores = os.cpu_count() or 1
    order = sorted(db.owned, key=lambda p: (p * 2654435761 + 12345) % (1 << 32))
    times = {p: oracle_bind.time_query_subset(db, query, [p], faithful=True, threads=cores)[0] for p in order}
    full = sum(times.values())
    sample, t = [], 0.0
    for p in order:
        sample.append(p)
        t += times[p]
        if t >= budget_s:
            break
    return full, sample, t, cores, times


def cpu_baseline_line(db, query, rows):
    """The in-line CPU baseline of the default run: a BOUNDED sample (10-30 s of CPU work) of the same query. The reference's
    Mutations loop (mutations.cpp:66-136) walks the positions in chunks of 300 over its thread pool and spends, per position
    and partition, O(|F_p|) contains() calls: its cost is (filtered rows) x (positions) -- measured at 1 M rows: 136-165 us per
    filtered row in every partition, and 20.7 s estimated from a third of the positions against 20.3-23.1 s measured. The sample
    cuts both ways -- about 30 % of the filtered rows (partitions in a fixed pseudo-random order) on the first 600 positions per
    host thread -- and is scaled by filtered rows x positions. `--impl reference` measures a full pass and prints per-partition
    times next to the filtered rows, so the estimator can be checked."""
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_bind
    cores = os.cpu_count() or 1
    flt = query["filterExpression"]
    filtered = {p: oracle_bind.filter_cardinality(db, flt, p) for p in db.owned}
    total_filtered = sum(filtered.values())
    order = sorted((p for p in db.owned if filtered[p]), key=lambda p: (p * 2654435761 + 12345) % (1 << 32))
    sample, seen = [], 0
    for p in order:
        sample.append(p)
        seen += filtered[p]
        if seen >= 0.3 * total_filtered:
            break
    positions = min(GENOME_LENGTH, 600 * max(cores, 1))      # two chunks of 300 positions per host thread: fewer leave threads idle and over-weigh the fixed costs
    t = oracle_bind.time_query_sample(db, query, sample, positions, threads=cores)
    full = t * (total_filtered / max(seen, 1)) * (GENOME_LENGTH / positions)
    fast_s = oracle_bind.time_query_subset(db, query, list(db.owned), faithful=False, threads=cores)[0]
    return {"value": 1.0 / full if full else None, "unit": UNIT, "cores": cores, "kind": "port",
            "best_cpu_algorithm": {"value": 1.0 / fast_s if fast_s else None, "unit": UNIT, "seconds": fast_s,
                                   "note": "NOT the reference's algorithm: the oracle's fast mode (difference array instead of the O(|F| L) contains() loop), whole query, all host threads"},
            "sample": f"oracle faithful mode (the reference's per-position loops, all {cores} host threads), rotation query 0 on {len(sample)} of the {len(order)} partitions the filter selects "
                      f"({seen} of {total_filtered} filtered rows) and positions 0..{positions} of {GENOME_LENGTH}: {t:.1f} s of CPU work, scaled by filtered rows x positions to {full:.1f} s for the whole {rows}-row query"}


def run_reference(args, rank, world):
    """The reference arm: the CPU restatement of the reference's own algorithm on the host cores (rank 0 only). A whole query
    takes minutes at 40 M rows, so a step is a bounded sample -- one selected partition on the first 300 positions per host
    thread -- and value = (share of the query the sample is) / (step time); the share comes from a short calibration pass over
    every selected partition. --full-pass measures the share with one pass of the whole query instead (minutes)."""
    if rank != 0:
        return
    import pysilo
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_bind
    t0 = time.time()
    db = pysilo.Database.synthetic(args.rows, args.seed)
    query = rotation(db)[0]
    cores = os.cpu_count() or 1
    filtered = {p: oracle_bind.filter_cardinality(db, query["filterExpression"], p) for p in db.owned}
    total_filtered = sum(filtered.values())
    selected = [p for p in db.owned if filtered[p]]
    # Calibration (seconds, not minutes): every selected partition on a short prefix of the positions gives its WEIGHT in the
    # query's CPU time. The per-partition cost is not a function of the filtered rows alone (measured full pass at 40 M rows,
    # profiles/r02_reference_arm_40m_full_pass.json: 36-64 us per filtered row depending on the filter's density), so the
    # share of a sample is measured, not derived; the cost is linear in the positions walked (same file: a partition on 16 %
    # of the positions took 16.3 % of its full time).
    calibration_positions = min(GENOME_LENGTH, 75 * cores)
    weight = {p: oracle_bind.time_query_sample(db, query, [p], calibration_positions, threads=cores) for p in selected}
    total_weight = sum(weight.values())
    # a step = the cheapest partition that still carries at least 5 % of the query, on the first 300 positions per host thread
    candidates = [p for p in selected if weight[p] >= 0.05 * total_weight] or selected
    sample = [min(candidates, key=lambda p: weight[p])]
    seen = sum(filtered[p] for p in sample)
    positions = min(GENOME_LENGTH, 300 * cores)
    fraction = (sum(weight[p] for p in sample) / max(total_weight, 1e-12)) * (positions / GENOME_LENGTH)
    fraction_source = f"measured weights of the selected partitions on {calibration_positions} positions x share of the positions"
    extra = {"calibration": [{"partition": p, "rows": db.partition_rows(p), "filtered_rows": filtered[p], "weight": round(weight[p] / total_weight, 5)} for p in selected]}
    if args.full_pass:
        full, _, _, _, per_partition = cpu_calibrate(db, query)
        t_sample = oracle_bind.time_query_sample(db, query, sample, positions, threads=cores)
        extra.update({"full_pass_s": full, "calibrated_fraction": fraction,
                      "per_partition": [{"partition": p, "rows": db.partition_rows(p), "filtered_rows": filtered[p], "cpu_s": round(per_partition[p], 3)} for p in db.owned]})
        fraction = t_sample / full
        fraction_source = "measured: sample time / time of one full pass over all partitions"
    setup_s = time.time() - t0

    def one_step():
        return oracle_bind.time_query_sample(db, query, sample, positions, threads=cores)
    for _ in range(min(args.warmup, 1)):            # a CPU arm needs no more than one warm-up pass
        one_step()
    times = [one_step() for _ in range(args.steps)]
    per_step = sum(times) / len(times)
    value = fraction / per_step                     # whole queries per second
    sample_text = (f"oracle faithful mode (the reference's per-position loops), all {cores} host threads, rotation query 0; each step = the query on {len(sample)} of the "
                   f"{sum(1 for p in db.owned if filtered[p])} partitions the filter selects ({seen} of {total_filtered} filtered rows), positions 0..{positions} of {GENOME_LENGTH} "
                   f"= {fraction:.5f} of the query ({fraction_source}); value = share / step time; the reference itself is not buildable here (DESIGN.md 2)")
    config = {"workload": workload_name(args.rows), "rows": args.rows, "partitions": db.n_partitions, "sample_fraction_of_query": fraction,
              "note": "queries/s of the host CPU: the same for any --gpus"}
    config.update(extra)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "strong" if args.gpus > 1 and not args.weak else "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": config,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample_text},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "setup_s": round(setup_s, 1)}
    print(json.dumps(line), flush=True)


def table_digest(np, table):
    valid = table[:VALID_ROWS * GENOME_LENGTH]
    return {"sum": int(valid.astype(np.uint64).sum()), "crc32": zlib.crc32(valid.tobytes())}


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    import pysilo
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    t0 = time.time()
    weak = world > 1 and args.weak
    units_per_step = world if weak else 1          # weak: every rank answers the query on a shard of its own
    slices = world if (world > 1 and not weak and args.shard == "rows") else 1
    db = pysilo.Database.synthetic(args.rows, args.seed + rank) if weak else pysilo.Database.synthetic(args.rows, args.seed, world=world, rank=rank, row_slices=slices)
    total_rows = args.rows * (world if weak else 1)
    t_generate = time.time() - t0
    engine = pysilo.Engine(db, local_rank, db.owned)
    t_ingest = time.time() - t0 - t_generate
    queries = rotation(db)
    texts = [json.dumps(q).encode() for q in queries]
    table_len = SIGMA * GENOME_LENGTH
    root = rank == 0

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def all_sum(values):
        if world == 1:
            return [int(v) for v in values]
        t = torch.tensor(values, dtype=torch.int64, device=device)
        dist.all_reduce(t)
        return [int(v) for v in t.cpu()]

    def all_max(values):
        if world == 1:
            return [float(v) for v in values]
        t = torch.tensor(values, dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.cpu()]

    # ---- before the communicator exists: every rank's LOCAL tables (valid-symbol rows), to check the reduced ones against ----
    local_sums, alg_bytes, k1_bytes = [], [], []
    for q in queries:
        p = pysilo.Prepared(engine, q, use_graph=False)
        p.run_device()
        local_sums.append(table_digest(np, p.table(0, table_len))["sum"])
        b = p.algorithmic_bytes(0)
        alg_bytes.append(b["alg_bytes"])
        k1_bytes.append(b["k1_bytes"])
        p.close()
    expected_sums = all_sum(local_sums)
    local_k1_bytes = list(k1_bytes)
    alg_bytes, k1_bytes = all_sum(alg_bytes), all_sum(k1_bytes)
    if world > 1:      # the library sums the count tables over the ranks itself (ncclReduce to rank 0 inside the query's submission)
        id_t = torch.zeros(128, dtype=torch.uint8, device=device)
        if root:
            id_t.copy_(torch.frombuffer(bytearray(pysilo.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(id_t, 0)
        pysilo.comm_init(engine, bytes(id_t.cpu().numpy().tobytes()), rank, world)
    prepared = [pysilo.Prepared(engine, q, use_graph=(world == 1)) for q in queries]

    def device_step(i):
        return prepared[i % N_ROTATION].run_device()      # synchronous; for N > 1 it includes the NCCL reduce of the table to rank 0

    def e2e_step(i):
        if world == 1:
            return engine.execute_text(texts[i % N_ROTATION])     # the JSON response text, as the reference's handler returns it
        return engine.execute_sharded(texts[i % N_ROTATION], root)

    warmup = max(args.warmup, N_ROTATION)
    for i in range(warmup):
        device_step(i)
    # ---- the reduced tables: rank 0's table of every query against the sum of the ranks' local tables ----
    digests = []
    for k in range(N_ROTATION):
        device_step(k)
        if root:
            d = table_digest(np, prepared[k].table(0, table_len))
            assert d["sum"] == expected_sums[k], f"query {k}: reduced table sums to {d['sum']}, the ranks' local tables to {expected_sums[k]}"
            digests.append(d)

    with ClockSampler(local_rank) as clocks:
        # ---- value: prepared queries, CUDA events around the timed steps, max over ranks ----
        barrier()
        launches0 = pysilo.launch_count(engine)
        pysilo.timer_begin(engine)
        result_rows = 0
        for i in range(args.steps):
            result_rows = device_step(i)
        device_ms = pysilo.timer_end(engine)
        barrier()
        launches = pysilo.launch_count(engine) - launches0
        # per query: CUDA-event time on the query's own stream, first kernel to last copy (median of 5 runs)
        gpu_ms = []
        for k in range(N_ROTATION):
            samples = []
            for _ in range(5):
                device_step(k)
                samples.append(prepared[k].elapsed_ms())
            gpu_ms.append(statistics.median(samples))
        gpu_ms = all_max(gpu_ms)
        # ---- throughput with several queries in flight: one host thread per prepared query ----
        throughput = None
        if world == 1 and args.in_flight > 1:
            q_threads = min(args.in_flight, N_ROTATION)
            per_thread = max(25, (16 * args.steps + q_threads - 1) // q_threads)

            def concurrent(work):
                start = threading.Barrier(q_threads + 1)
                def loop(k):
                    work(k)
                    start.wait()
                    for _ in range(per_thread):
                        work(k)
                threads = [threading.Thread(target=loop, args=(k,)) for k in range(q_threads)]
                for t in threads:
                    t.start()
                torch.cuda.synchronize()
                start.wait()
                pysilo.timer_begin(engine)
                w = time.perf_counter()
                for t in threads:
                    t.join()
                ms = pysilo.timer_end(engine)                  # every call is synchronous: all GPU work is done here
                return q_threads * per_thread / (ms * 1e-3), q_threads * per_thread / (time.perf_counter() - w)

            dev_qps, _ = concurrent(lambda k: prepared[k].run_device())
            engine.set_plan_cache(False)
            _, e2e_qps = concurrent(lambda k: engine.execute_text(texts[k]))
            throughput = {"queries_in_flight": q_threads, "queries": q_threads * per_thread, "value": dev_qps, "e2e_value": e2e_qps, "unit": UNIT,
                          "how": "one host thread per rotation query calling the same entry points concurrently; prepared replay timed with CUDA events, e2e (plan cache off) by wall clock"}
        # ---- the container scan timed alone (kernels of a query serialised on one stream): the roofline line, rotation query 0 ----
        engine.set_plan_cache(False)
        pysilo.set_overlap(engine, False)
        for _ in range(3):
            e2e_step(0)
        pysilo.k1_accumulated(engine, reset=True)
        for _ in range(max(args.steps // 2, 10)):
            e2e_step(0)
        k1_alone_ms, k1_alone_launches = pysilo.k1_accumulated(engine, reset=True)
        pysilo.set_overlap(engine, True)
        for _ in range(3):
            e2e_step(0)
        pysilo.k1_accumulated(engine, reset=True)
        for _ in range(max(args.steps // 2, 10)):
            e2e_step(0)
        k1_ms, k1_launches = pysilo.k1_accumulated(engine, reset=True)
        barrier()
        # ---- e2e: public call with host buffers, plan cache off, wall clock bracketed by barrier + synchronize ----
        for i in range(N_ROTATION):
            e2e_step(i)
        h2d0, d2h0 = pysilo.transfer_bytes(engine)
        barrier()
        w0 = time.perf_counter()
        for i in range(args.steps):
            result = e2e_step(i)
        barrier()
        e2e_s = time.perf_counter() - w0
        h2d1, d2h1 = pysilo.transfer_bytes(engine)
        # ---- the same with the engine's plan cache (query text -> prepared query; a CUDA graph at N = 1) ----
        e2e_cached_s = None
        if world == 1:
            engine.set_plan_cache(True)
            for i in range(3 * N_ROTATION):
                e2e_step(i)
            barrier()
            w0 = time.perf_counter()
            for i in range(args.steps):
                result = e2e_step(i)
            barrier()
            e2e_cached_s = time.perf_counter() - w0
    device_ms, e2e_s = all_max([device_ms, e2e_s])
    # roofline line: the rank whose container scan of query 0 has the most bytes (a rank may own no selected partition at all)
    roofline_rank = 0
    mine = [float(local_k1_bytes[0]), k1_ms, float(k1_launches), k1_alone_ms, float(k1_alone_launches)]
    if world > 1:
        t = torch.tensor(mine, dtype=torch.float64, device=device)
        every = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(every, t)
        roofline_rank = max(range(world), key=lambda r: float(every[r][0]))
        mine = [float(x) for x in every[roofline_rank].cpu()]
    scan_bytes, k1_ms, k1_launches, k1_alone_ms, k1_alone_launches = mine

    extra = None
    if args.lineage and world == 1:
        extra = lineage_queries(pysilo, db, engine)

    if root:
        peak, peak_source = measured_peak()
        k1_instep_ms = k1_ms / max(k1_launches, 1)          # with the missing-symbol kernel beside it
        k1_avg_ms = k1_alone_ms / max(k1_alone_launches, 1)  # the same launches with nothing beside them
        achieved = scan_bytes / (k1_avg_ms * 1e-3) / 1e9 if k1_avg_ms > 0 else 0.0
        traffic = None
        traffic_file = ROOT / "profiles" / "k1_traffic.json"
        if traffic_file.exists() and world == 1:
            traffic = json.loads(traffic_file.read_text()).get(str(args.rows))
        stats = pysilo.pool_stats(engine)
        ms_per_step = device_ms / args.steps
        mean_alg = sum(alg_bytes) / len(alg_bytes)
        mean_gpu_ms = sum(gpu_ms) / len(gpu_ms)
        line = {
            "metric": METRIC, "value": units_per_step * args.steps / (device_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak" if weak or world == 1 else "strong", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": workload_name(args.rows), "rows": total_rows, "partitions": db.n_partitions * (world if weak else 1), "partitions_on_rank0": len(db.owned),
                       "l2": "inputs larger than L2 (container pool streamed once per query; 8 different queries in rotation)",
                       "parallelism": (f"{world} replicated shard(s) of {args.rows} sequences (seed + rank), one per GPU" if weak else
                                       f"one database of {args.rows} sequences, " + (f"every lineage partition cut into {world} DatabasePartitions of consecutive rows, one per GPU" if slices > 1 else f"whole lineage DatabasePartitions assigned to {world} GPU(s) by rows (LPT)") +
                                       ", NCCL reduce of the valid-symbol rows of the count table inside the query's submission"),
                       "shard": args.shard if world > 1 and not weak else None},
            "e2e": {"value": units_per_step * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": (h2d1 - h2d0) // args.steps + sum(len(t) for t in texts) // len(texts),
                    "d2h_bytes_per_step": (d2h1 - d2h0) // args.steps, "ms_per_step": e2e_s * 1e3 / args.steps, "plan_cache": "off",
                    "cached": {"value": args.steps / e2e_cached_s, "ms_per_step": e2e_cached_s * 1e3 / args.steps, "note": "query text -> prepared query (CUDA graph) in the engine's plan cache"} if e2e_cached_s else None},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "k1_scan", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "peak_source": peak_source, "rank": roofline_rank, "query": 0, "algorithmic_bytes_per_launch": scan_bytes, "avg_launch_ms": k1_avg_ms,
                         "avg_launch_ms_overlapped": k1_instep_ms, "achieved_overlapped": scan_bytes / (k1_instep_ms * 1e-3) / 1e9 if k1_instep_ms > 0 else 0.0,
                         "timing": "CUDA events around the launch on the library stream (one-shot entry point); avg_launch_ms with the query's kernels serialised (sb_set_overlap 0), _overlapped with the missing-symbol kernel beside it",
                         "share_of_query": k1_instep_ms / gpu_ms[0] if gpu_ms[0] else None},
            "query_gpu_ms": gpu_ms, "query_algorithmic_bytes": alg_bytes, "query_k1_bytes": k1_bytes,
            "scan_gb_per_s": mean_alg / (mean_gpu_ms * 1e-3) / 1e9 if mean_gpu_ms else None,
            "scan_note": "mean algorithmic bytes of the rotation's queries / mean CUDA-event time of a query on its own stream (first kernel to last copy; N > 1: slowest rank)",
            "table_digests": digests, "table_check": "rank 0's (reduced) table of every query sums to the sum of the ranks' local tables over the valid-symbol rows",
            "throughput": throughput,
            "clocks": clocks.summary(),
            "pool": {k: stats[k] for k in ("array_containers", "run_containers", "bitset_containers", "payload_bytes", "directory_bytes", "missing_bytes", "full_table_bytes", "n_work_items")},
            "result_rows": int(result_rows),
            "other_queries": extra,
            "setup_s": {"generate": round(t_generate, 1), "ingest": round(t_ingest, 1)},
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_line(db, queries[0], args.rows)
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)
    for p in prepared:
        p.close()
    if world > 1:
        pysilo.comm_free(engine)
        dist.barrier()
        dist.destroy_process_group()


def lineage_queries(pysilo, db, engine):
    """C5's other two filters: a lineage with sublineages (full / partial / empty partitions) and True (all full)"""
    names = sorted({".".join(n.split(".")[:k]) for n in db.lineage_names() for k in range(1, n.count(".") + 2)}, key=lambda n: (n.count("."), n))
    out = []
    for label, flt in [(f"PangoLineage {names[len(names) // 3]}*", {"type": "PangoLineage", "column": "pangoLineage", "value": names[len(names) // 3], "includeSublineages": True}),
                       (f"PangoLineage {names[2]}*", {"type": "PangoLineage", "column": "pangoLineage", "value": names[2], "includeSublineages": True}), ("True", {"type": "True"})]:
        q = {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": flt}
        p = pysilo.Prepared(engine, q)
        samples = []
        for _ in range(8):
            rows = p.run_device()
            samples.append(p.elapsed_ms())
        text = json.dumps(q).encode()
        engine.set_plan_cache(False)
        for _ in range(3):
            engine.execute_text(text)
        w = time.perf_counter()
        for _ in range(10):
            engine.execute_text(text)
        out.append({"filter": label, "gpu_ms": statistics.median(samples), "e2e_ms": (time.perf_counter() - w) * 100, "result_rows": rows, "algorithmic_bytes": p.algorithmic_bytes(0)["alg_bytes"],
                    "note": "full partitions are table hits (precomputed at load): their algorithmic bytes are not read"})
        p.close()
    return out


if __name__ == "__main__":
    main()
