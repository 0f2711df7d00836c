#!/usr/bin/env python
"""Profiling harness (diagnostics, not part of the product path): the bench workload with the timed queries bracketed by
cudaProfilerStart / cudaProfilerStop, for `ncu --profile-from-start off ...`. Without ncu it prints per-query CUDA-event
times and the container scan's time alone / overlapped.

    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
        python tools/profile_query.py --rows 40000000 --queries 0,1 --runs 1
"""
import argparse
import json
import statistics
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "lapis-silo_b200"))
sys.path.insert(0, str(ROOT))


def sweep_variants(args):
    """every variant in ONE process (the 40 M-row database takes two minutes to build): container scan alone (one stream) and
    the whole query (two streams), per rotation query; the count tables of all variants must be identical"""
    import ctypes
    import zlib
    import pysilo
    import bench
    db = pysilo.Database.synthetic(args.rows, args.seed)
    engine = pysilo.Engine(db, 0, db.owned)
    lib = pysilo._sb()
    lib.sb_debug_set_k1_variant.argtypes = [ctypes.c_void_p, ctypes.c_uint32]
    queries = bench.rotation(db)
    picks = [int(k) for k in args.queries.split(",")]
    texts = {k: json.dumps(queries[k]).encode() for k in picks}
    engine.set_plan_cache(False)
    table_len = 16 * 29903
    reference = {}
    for variant in [int(v) for v in args.variants.split(",")]:
        assert lib.sb_debug_set_k1_variant(engine.sb, variant) == 0
        line = {"variant": variant, "rows": args.rows}
        for k in picks:
            p = pysilo.Prepared(engine, queries[k], use_graph=False)
            p.run_device()
            crc = zlib.crc32(p.table(0, table_len).tobytes())
            reference.setdefault(k, crc)
            line[f"q{k}_table_ok"] = crc == reference[k]
            for _ in range(args.warmup):
                p.run_device()
            samples = []
            for _ in range(max(args.runs, 5)):
                p.run_device()
                samples.append(p.elapsed_ms())
            line[f"q{k}_query_ms"] = statistics.median(samples)
            pysilo.set_overlap(engine, False)
            for _ in range(args.warmup):
                engine.execute_text(texts[k])
            pysilo.k1_accumulated(engine, reset=True)
            for _ in range(max(args.runs, 5)):
                engine.execute_text(texts[k])
            ms, n = pysilo.k1_accumulated(engine, reset=True)
            pysilo.set_overlap(engine, True)
            k1_bytes = p.algorithmic_bytes(0)["k1_bytes"]
            line[f"q{k}_k1_ms"] = ms / max(n, 1)
            line[f"q{k}_k1_gbs"] = k1_bytes / (ms / max(n, 1) * 1e-3) / 1e9 if ms else None
            p.close()
        print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=40_000_000)
    ap.add_argument("--seed", type=lambda s: int(s, 0), default=0x51100005)
    ap.add_argument("--queries", default="0", help="rotation members to run inside the profiled region")
    ap.add_argument("--runs", type=int, default=1)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--serial", action="store_true", help="one stream (sb_set_overlap 0)")
    ap.add_argument("--graph", action="store_true")
    ap.add_argument("--one-shot", action="store_true", help="sb_query_mutations through the engine instead of a prepared query")
    ap.add_argument("--variants", default="", help="comma list of container-scan variants (sb_debug_set_k1_variant) to time side by side; no profiler region")
    args = ap.parse_args()
    if args.variants:
        return sweep_variants(args)
    import torch
    import pysilo
    import bench
    db = pysilo.Database.synthetic(args.rows, args.seed)
    engine = pysilo.Engine(db, 0, db.owned)
    if args.serial:
        pysilo.set_overlap(engine, False)
    queries = bench.rotation(db)
    picks = [int(k) for k in args.queries.split(",")]
    prepared = {k: pysilo.Prepared(engine, queries[k], use_graph=args.graph) for k in picks}
    texts = {k: json.dumps(queries[k]).encode() for k in picks}
    engine.set_plan_cache(False)

    def run(k):
        if args.one_shot:
            engine.execute_text(texts[k])
            return 0.0
        prepared[k].run_device()
        return prepared[k].elapsed_ms()

    for _ in range(args.warmup):
        for k in picks:
            run(k)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    times = {k: [] for k in picks}
    pysilo.k1_accumulated(engine, reset=True)
    for _ in range(args.runs):
        for k in picks:
            times[k].append(run(k))
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    k1_ms, k1_n = pysilo.k1_accumulated(engine, reset=True)
    out = {"rows": args.rows, "serial": args.serial, "graph": args.graph,
           "query_gpu_ms": {k: statistics.median(v) for k, v in times.items()},
           "k1_bytes": {k: prepared[k].algorithmic_bytes(0)["k1_bytes"] for k in picks},
           "alg_bytes": {k: prepared[k].algorithmic_bytes(0)["alg_bytes"] for k in picks},
           "k1_ms_one_shot": k1_ms / k1_n if k1_n else None}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
