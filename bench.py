#!/usr/bin/env python
"""bench.py — Mutations-action queries/s on synthetic SARS-CoV-2-shaped data (BASELINE.json).

One "step" = one nucleotide Mutations query (minProportion 0.05) under the N-Of filter of BASELINE config
C3 (3 of 10 clade-defining NucleotideEquals) over one shard of 10 M sequences x 29,903 nt. The path shards by
DatabasePartition: with N GPUs every rank holds the DatabasePartitions of its own 10 M-sequence shard (weak
scaling: an N x 10 M-sequence database; the shards are statistically identical, generated from the same seed), a
step runs the query on every shard and sums the per-rank count tables with an NCCL reduce, and `value` counts one
query per shard: N x steps / time (at N = 1 plainly queries/s). `--strong` instead shards ONE database of
`--rows` sequences over the ranks (the 40 M-sequence configuration C5: `--rows 40000000 --strong`).

    value  queries/s with the index resident in HBM and the query compiled once: filter kernels + count
           kernels (+ NCCL reduce), timed with CUDA events on the library's stream, max over ranks
    e2e    queries/s through the public call a user makes: query JSON in -> result JSON out, i.e.
           parse, FilterExpression -> Operator compile, plan upload (H2D), kernels, count table download
           (D2H), thresholds / proportions on the host
    --impl reference   the CPU restatement of the reference's algorithm (oracle/, all host threads) on a
           bounded sample of the same workload (the reference itself cannot be built here)
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT / "lapis-silo_b200"))

GENOME_LENGTH = 29903
SIGMA = 16
METRIC = "mutations_queries_per_s"
UNIT = "queries/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=int(os.environ.get("SILO_BENCH_ROWS", 10_000_000)))
    ap.add_argument("--seed", type=lambda s: int(s, 0), default=0x51100003)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--strong", action="store_true", help="N > 1: shard one database of --rows sequences over the ranks instead of one shard of --rows per rank")
    ap.add_argument("--in-flight", type=int, default=8, help="queries in flight of the extra throughput measurement (N = 1 only; 0 = skip)")
    return ap.parse_args()


def c3_query(db):
    clade = db.clade_mutations()
    picks = [clade[(7 * k) % len(clade)] for k in range(10)]
    return {"action": {"type": "Mutations", "minProportion": 0.05},
            "filterExpression": {"type": "N-Of", "numberOfMatchers": 3, "matchExactly": False,
                                 "children": [{"type": "NucleotideEquals", "position": p + 1, "symbol": s} for p, s in picks]}}


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
            self.stop.wait(0.2)

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


def cpu_sample(db, query, total_rows, budget_s=12.0):
    """Oracle (CPU restatement of the reference's algorithm, faithful mode, all host threads) on a bounded sample of the
    workload: the query restricted to one partition at a time, partitions taken in a fixed pseudo-random order (the cost
    of a partition depends on how many of its rows the filter selects, so neither the smallest nor the largest are
    representative), until about `budget_s` seconds of CPU work are spent or every partition was visited; scaled to the
    whole database by rows: value = (sample_rows / total_rows) / seconds  [queries/s]."""
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_bind
    cores = os.cpu_count() or 1
    order = sorted(db.owned, key=lambda p: (p * 2654435761 + 12345) % (1 << 32))
    sample, rows, seconds = [], 0, 0.0
    for p in order:
        t, _ = oracle_bind.time_query_subset(db, query, [p], faithful=True, threads=cores)
        sample.append(p)
        rows += db.partition_rows(p)
        seconds += t
        if seconds >= budget_s:
            break
    what = "every partition" if rows == total_rows else f"{len(sample)} of {len(order)} partitions = {rows} of {total_rows} rows, scaled by rows"
    return {"value": (rows / total_rows) / seconds, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"oracle faithful mode, one partition at a time in a fixed pseudo-random order, {what}, {seconds:.1f} s of CPU work"}, sample


def run_reference(args, rank, world):
    """The reference arm: the CPU restatement of the reference's own algorithm on the host cores."""
    if rank != 0:
        return
    import pysilo
    t0 = time.time()
    db = pysilo.Database.synthetic(args.rows, args.seed)
    query = c3_query(db)
    base, sample = cpu_sample(db, query, args.rows, budget_s=4.0)     # the per-step sample: about 4 s of CPU work
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_bind
    cores = os.cpu_count() or 1
    sample_rows = sum(db.partition_rows(p) for p in sample)
    def one_step():
        return sum(oracle_bind.time_query_subset(db, query, [p], faithful=True, threads=cores)[0] for p in sample)
    for _ in range(min(args.warmup, 1)):       # a CPU arm needs no more than one warm-up pass
        one_step()
    times = [one_step() for _ in range(args.steps)]
    per_step = sum(times) / len(times)
    value = (sample_rows / args.rows) / per_step
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": per_step * 1e3 * args.rows / sample_rows, "higher_is_better": True, "scaling": "strong" if args.strong and args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "u32",
            "data": "synthetic", "config": {"workload": f"C3: {args.rows} synthetic sequences x {GENOME_LENGTH} nt per shard, nucleotide Mutations under N-Of(3 of 10)", "rows": args.rows,
                                            "note": "shard-queries/s of the host CPU: the same for any --gpus (the CPU answers the shards one after the other)"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": base["sample"] + " per step; the reference itself is not buildable here (DESIGN.md 2)"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "setup_s": round(time.time() - t0, 1)}
    print(json.dumps(line), flush=True)


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
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    t0 = time.time()
    weak = world > 1 and not args.strong
    units_per_step = world if weak else 1          # shard-queries per step (weak scaling: every rank answers the query on its own shard)
    db = pysilo.Database.synthetic(args.rows, args.seed) if weak else pysilo.Database.synthetic(args.rows, args.seed, world=world, rank=rank)
    total_rows = args.rows * (world if weak else 1)
    t_generate = time.time() - t0
    engine = pysilo.Engine(db, local_rank, db.owned)
    t_ingest = time.time() - t0 - t_generate
    query = c3_query(db)
    query_json = json.dumps(query)
    query_bytes = query_json.encode()
    table_len = SIGMA * GENOME_LENGTH
    counts = torch.zeros(table_len, dtype=torch.int32, device=f"cuda:{local_rank}")
    prepared = pysilo.Prepared(engine, query)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    if world > 1:      # the library sums the count tables over the ranks itself (ncclReduce to rank 0 on the stream of the count kernels)
        id_t = torch.zeros(128, dtype=torch.uint8, device=counts.device)
        if rank == 0:
            id_t.copy_(torch.frombuffer(bytearray(pysilo.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(id_t, 0)
        pysilo.comm_init(engine, bytes(id_t.cpu().numpy().tobytes()), rank, world)

    def device_step():
        prepared.run_device(0, counts.data_ptr(), table_len)      # synchronous; for N > 1 it ends with the NCCL reduce of the table to rank 0

    def e2e_step():
        if world == 1:
            return engine.execute_text(query_bytes)     # the JSON response text, as the reference's handler returns it
        q = pysilo.DistributedQuery(engine, query_json)
        q.local_table_device(0, counts.data_ptr(), table_len)      # local filter + count table, NCCL reduce to rank 0
        result = None
        if rank == 0:
            q.set_table_device(0, counts.data_ptr())              # rows are selected on the device; only they come back
            result = q.finish_text()                               # (a table of zeros -- no rank produced anything -- selects no row)
        q.close()
        return result

    for _ in range(max(args.warmup, 3)):
        device_step()
        e2e_step()
    info = prepared.run_device(0, counts.data_ptr(), table_len, want_bytes=True)
    local_k1_bytes = info["k1_bytes"]
    if world > 1:
        bytes_t = torch.tensor([info["alg_bytes"], info["k1_bytes"]], dtype=torch.int64, device=counts.device)
        dist.all_reduce(bytes_t)
        info["alg_bytes"], info["k1_bytes"] = int(bytes_t[0]), int(bytes_t[1])

    with ClockSampler(local_rank) as clocks:
        # ---- value: device-resident steps, CUDA events on the library's stream, max over ranks ----
        barrier()
        pysilo.k1_accumulated(engine, reset=True)
        launches0 = pysilo.launch_count(engine)
        pysilo.timer_begin(engine)
        for _ in range(args.steps):
            device_step()
        device_ms = pysilo.timer_end(engine)
        barrier()
        launches = pysilo.launch_count(engine) - launches0
        k1_ms, k1_launches = pysilo.k1_accumulated(engine, reset=True)
        # ---- throughput with several queries in flight: host threads, one library lane (streams + scratch) each ----
        throughput = None
        if world == 1 and args.in_flight > 1:
            q_threads = args.in_flight
            per_thread = max(25, (16 * args.steps + q_threads - 1) // q_threads)
            tables = [torch.zeros(table_len, dtype=torch.int32, device=counts.device) for _ in range(q_threads)]

            def concurrent(work):
                start = threading.Barrier(q_threads + 1)
                def loop(k):
                    work(k)                                   # binds the thread to its lane, grows the lane's buffers
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

            dev_qps, _ = concurrent(lambda k: prepared.run_device(0, tables[k].data_ptr(), table_len))
            _, e2e_qps = concurrent(lambda k: engine.execute_text(query_bytes))
            throughput = {"queries_in_flight": q_threads, "queries": q_threads * per_thread, "value": dev_qps, "e2e_value": e2e_qps, "unit": UNIT,
                          "how": "host threads calling the same entry points concurrently; device-resident replay timed with CUDA events, e2e by wall clock"}
        # ---- the container scan timed alone (kernels of a query serialised on one stream): the roofline line ----
        pysilo.set_overlap(engine, False)
        for _ in range(2):
            device_step()
        pysilo.k1_accumulated(engine, reset=True)
        for _ in range(args.steps):
            device_step()
        k1_alone_ms, k1_alone_launches = pysilo.k1_accumulated(engine, reset=True)
        pysilo.set_overlap(engine, True)
        barrier()
        # ---- e2e: public API with host buffers, wall clock bracketed by barrier + synchronize ----
        h2d0, d2h0 = pysilo.transfer_bytes(engine)
        barrier()
        w0 = time.perf_counter()
        for _ in range(args.steps):
            result = e2e_step()
        barrier()
        e2e_s = time.perf_counter() - w0
        h2d1, d2h1 = pysilo.transfer_bytes(engine)
    if world > 1:
        t = torch.tensor([device_ms, e2e_s], dtype=torch.float64, device=counts.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        device_ms, e2e_s = float(t[0]), float(t[1])
        # roofline line: the rank whose container scan has the most bytes (rank 0 may own no selected partition at all)
        mine = torch.tensor([float(local_k1_bytes), k1_ms, float(k1_launches), k1_alone_ms, float(k1_alone_launches)], dtype=torch.float64, device=counts.device)
        every = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        best = max(range(world), key=lambda r: float(every[r][0]))
        local_k1_bytes, k1_ms, k1_launches, k1_alone_ms, k1_alone_launches = (float(x) for x in every[best])
        local_k1_bytes, roofline_rank = int(local_k1_bytes), best
    else:
        roofline_rank = 0

    if rank == 0:
        peak, peak_source = measured_peak()
        k1_instep_ms = k1_ms / max(k1_launches, 1)          # inside the timed steps: the missing-symbol kernel runs beside it
        k1_avg_ms = k1_alone_ms / max(k1_alone_launches, 1)  # the same launches with nothing beside them
        achieved = local_k1_bytes / (k1_avg_ms * 1e-3) / 1e9 if k1_avg_ms > 0 else 0.0
        traffic = None
        traffic_file = ROOT / "profiles" / "k1_traffic.json"
        if traffic_file.exists():
            traffic = json.loads(traffic_file.read_text()).get(str(args.rows)) if world == 1 else None
        stats = pysilo.pool_stats(engine)
        d2h_extra = 0
        line = {
            "metric": METRIC, "value": units_per_step * args.steps / (device_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": device_ms / args.steps, "higher_is_better": True, "scaling": "weak" if weak or world == 1 else "strong", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": f"C3: {args.rows} synthetic sequences x {GENOME_LENGTH} nt per shard, nucleotide Mutations under N-Of(3 of 10)", "rows": total_rows,
                       "rows_per_rank": args.rows if weak or world == 1 else None, "shard_queries_per_step": units_per_step,
                       "partitions": db.n_partitions * (world if weak else 1), "partitions_on_rank0": len(db.owned), "l2": "inputs larger than L2 (container pool streamed once per query)",
                       "parallelism": (f"{world} shard(s) of {args.rows} sequences, one per GPU (DatabasePartitions are the unit of sharding); a step = the query on every shard + NCCL reduce "
                                       f"of the count tables; value = shard-queries/s" if weak or world == 1 else f"one database of {args.rows} sequences, partitions sharded over {world} GPUs, NCCL reduce of the count table")},
            "e2e": {"value": units_per_step * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": (h2d1 - h2d0) // args.steps + len(query_json),
                    "d2h_bytes_per_step": (d2h1 - d2h0) // args.steps + d2h_extra, "ms_per_step": e2e_s * 1e3 / args.steps},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "k1_and_popcount", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "peak_source": peak_source, "rank": roofline_rank, "algorithmic_bytes_per_launch": local_k1_bytes, "avg_launch_ms": k1_avg_ms,
                         "avg_launch_ms_overlapped": k1_instep_ms, "achieved_overlapped": local_k1_bytes / (k1_instep_ms * 1e-3) / 1e9 if k1_instep_ms > 0 else 0.0,
                         "timing": "CUDA events around the launch on the library stream; avg_launch_ms with the query's kernels serialised (sb_set_overlap 0), _overlapped as they run in the timed steps",
                         "share_of_step": k1_instep_ms / (device_ms / args.steps) if device_ms else None},
            "throughput": throughput,
            "scan_gb_per_s": info["alg_bytes"] / (device_ms / args.steps * 1e-3) / 1e9,
            "algorithmic_bytes_per_query": info["alg_bytes"],
            "clocks": clocks.summary(),
            "pool": {k: stats[k] for k in ("array_containers", "run_containers", "bitset_containers", "payload_bytes", "directory_bytes", "missing_bytes", "full_table_bytes", "n_work_items")},
            "result_rows": (result.count('"mutation"') if isinstance(result, str) else len(result)) if result is not None else None,
            "setup_s": {"generate": round(t_generate, 1), "ingest": round(t_ingest, 1)},
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"], _ = cpu_sample(db, query, args.rows)
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)
    prepared.close()
    if world > 1:
        pysilo.comm_free(engine)
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
