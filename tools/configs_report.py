#!/usr/bin/env python
"""Measures the BASELINE configurations C1-C5 on one B200 through the public call (query JSON in ->
response JSON out) and, for single-sequence Mutations queries, device-resident as bench.py does; one JSON line per query.

    python tools/configs_report.py [--only C1,C2] [--c4-rows 10000000] [--c5-rows 40000000] > gpurun_out/configs.jsonl

Per query: rows N, partitions P, pool bytes by container kind, algorithmic bytes of the query (Mutations), median / p99
latency and queries/s end to end (>= 20 timed calls after 3 warm-ups), device-resident step time, result rows.
Parity is the job of tests/ (the same query shapes at sizes the oracle finishes, C1 itself in test_config_c1_thousand_sequences);
this tool only measures and never touches oracle/.
"""
import argparse
import json
import statistics
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "lapis-silo_b200"))
sys.path.insert(0, str(ROOT))
GOLDEN = ROOT / "tests" / "golden"


def measure(pysilo, torch, engine, db, name, query, rows, device_table=None, calls=30):
    text = json.dumps(query).encode()
    for _ in range(3):
        body = engine.execute_text(text)
    lat = []
    for _ in range(calls):
        t0 = time.perf_counter()
        body = engine.execute_text(text)
        lat.append((time.perf_counter() - t0) * 1e3)
    lat.sort()
    stats = pysilo.pool_stats(engine)
    line = {"config": name, "rows": rows, "partitions": db.n_partitions, "query": query,
            "e2e_ms_median": statistics.median(lat), "e2e_ms_p99": lat[min(len(lat) - 1, int(0.99 * len(lat)))], "e2e_queries_per_s": 1e3 / statistics.median(lat),
            "result_rows": len(json.loads(body)["queryResult"]), "response_bytes": len(body),
            "pool": {k: stats[k] for k in ("array_containers", "run_containers", "bitset_containers", "payload_bytes", "missing_bytes", "total_device_bytes")}}
    if device_table:       # device-resident replay of the prepared query (one CUDA graph launch), CUDA events on its stream
        prepared = pysilo.Prepared(engine, query)
        for _ in range(3):
            prepared.run_device()
        samples = []
        torch.cuda.profiler.start()      # `ncu --profile-from-start off` sees the replays only
        for _ in range(calls):
            prepared.run_device()
            samples.append(prepared.elapsed_ms())
        torch.cuda.profiler.stop()
        ms = statistics.median(samples)
        info = prepared.algorithmic_bytes(device_table[0])
        line.update({"device_ms_per_query": ms, "device_queries_per_s": 1e3 / ms, "algorithmic_bytes": info["alg_bytes"], "k1_algorithmic_bytes": info["k1_bytes"],
                     "scan_gb_per_s": info["alg_bytes"] / ms / 1e6 if info["k1_bytes"] else None})      # no scan rate for pure table hits (full partitions: precomputed tables)
        prepared.close()
    print(json.dumps(line), flush=True)
    return body


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="C1,C2,C3,C4,C5")
    ap.add_argument("--c3-rows", type=int, default=10_000_000)
    ap.add_argument("--c4-rows", type=int, default=10_000_000)
    ap.add_argument("--c5-rows", type=int, default=40_000_000)
    args = ap.parse_args()
    only = set(args.only.split(","))
    import torch
    import pysilo
    import bench
    L = 29903
    if "C1" in only:       # the reference's own 1,000-sequence dataset, Mutations under date + lineage
        db = pysilo.Database.from_directory(GOLDEN / "exampleDataset1000Sequences")
        engine = pysilo.Engine(db, 0)
        q = {"action": {"type": "Mutations", "minProportion": 0.05},
             "filterExpression": {"type": "And", "children": [{"type": "DateBetween", "column": "date", "from": "2020-01-01", "to": "2021-12-31"},
                                                              {"type": "PangoLineage", "column": "pangoLineage", "value": "B.1", "includeSublineages": True}]}}
        measure(pysilo, torch, engine, db, "C1", q, 1000, (0, 16 * L))
        engine.close()
    if "C2" in only:       # 1 M sequences, Aggregated under And[Or[...], ...] with metadata predicates
        db = pysilo.Database.synthetic(1_000_000, 0x51100002)
        engine = pysilo.Engine(db, 0)
        c = db.clade_mutations()
        ne = lambda k: {"type": "NucleotideEquals", "position": c[k][0] + 1, "symbol": c[k][1]}
        q = {"action": {"type": "Aggregated"},
             "filterExpression": {"type": "And", "children": [{"type": "Or", "children": [ne(0), ne(1)]}, ne(2),
                                                              {"type": "Not", "child": {"type": "NucleotideEquals", "position": c[3][0] + 1, "symbol": "-"}},
                                                              {"type": "DateBetween", "column": "date", "from": "2021-01-01", "to": "2021-12-31"},
                                                              {"type": "StringEquals", "column": "region", "value": "Europe"},
                                                              {"type": "IntBetween", "column": "age", "from": 20, "to": 60}]}}
        measure(pysilo, torch, engine, db, "C2", q, 1_000_000)
        engine.close()
    if "C3" in only:       # bench.py's query
        db = pysilo.Database.synthetic(args.c3_rows, 0x51100003)
        engine = pysilo.Engine(db, 0)
        measure(pysilo, torch, engine, db, "C3", bench.c3_query(db), args.c3_rows, (0, 16 * L))
        engine.close()
    if "C4" in only:       # amino-acid Mutations over all genes under HasAminoAcidMutation / HasNucleotideMutation
        db = pysilo.Database.synthetic(args.c4_rows, 0x51100004, with_amino_acids=True)
        engine = pysilo.Engine(db, 0)
        for label, f in (("C4 (HasAminoAcidMutation S:501)", {"type": "HasAminoAcidMutation", "sequenceName": "S", "position": 501}),
                         ("C4 (HasNucleotideMutation 23063)", {"type": "HasNucleotideMutation", "position": 23063})):
            measure(pysilo, torch, engine, db, label, {"action": {"type": "AminoAcidMutations", "minProportion": 0.05}, "filterExpression": f}, args.c4_rows, (0, 0))
        engine.close()
    if "C5" in only:       # 40 M sequences: Mutations under a top-level clade with sublineages, and under True
        db = pysilo.Database.synthetic(args.c5_rows, 0x51100005)
        engine = pysilo.Engine(db, 0)
        clade = sorted(n for n in db.lineage_names() if n.count(".") == 2)[0]
        for label, f in ((f"C5 (PangoLineage {clade}*)", {"type": "PangoLineage", "column": "pangoLineage", "value": clade, "includeSublineages": True}), ("C5 (True)", {"type": "True"})):
            measure(pysilo, torch, engine, db, label, {"action": {"type": "Mutations", "minProportion": 0.05}, "filterExpression": f}, args.c5_rows, (0, 16 * L))
        engine.close()


if __name__ == "__main__":
    main()
