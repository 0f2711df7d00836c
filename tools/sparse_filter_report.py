#!/usr/bin/env python
"""Evidence for the dense filter design on SPARSE filters (diagnostics, not part of the product path).

The filter stage evaluates a plan per (partition, 65,536-row container key) and writes a dense result bitset; the reference
intersects two array containers by galloping (intersection.cpp:62-74) and never touches a row that is in neither. This
report times Aggregated counts under filters whose leaves are private mutations (a few hundred rows out of 40 M) and puts the
bytes the reference's algorithm would read (SURVEY 8d B_filter: 2 bytes per value + 8 per container) beside the time.

    python tools/sparse_filter_report.py --rows 40000000 > gpurun_out/sparse_filters.json
"""
import argparse
import json
import statistics
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "lapis-silo_b200"))
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=40_000_000)
    ap.add_argument("--seed", type=lambda s: int(s, 0), default=0x51100005)
    args = ap.parse_args()
    import pysilo
    db = pysilo.Database.synthetic(args.rows, args.seed)
    engine = pysilo.Engine(db, 0, db.owned)
    clade = {p for p, _ in db.clade_mutations()}
    n_keys = sum((db.partition_rows(p) + 65535) // 65536 for p in db.owned)

    def count(flt):
        return engine.execute({"action": {"type": "Aggregated"}, "filterExpression": flt})[0]["count"]

    def timed(flt):
        p = pysilo.Prepared(engine, {"action": {"type": "Aggregated"}, "filterExpression": flt}, use_graph=True)
        for _ in range(5):
            n = p.run_device()
        samples = []
        for _ in range(20):
            p.run_device()
            samples.append(p.elapsed_ms())
        p.close()
        return n, statistics.median(samples)

    leaves = []
    for pos in range(1000, 29000, 1777):
        if pos in clade:
            continue
        for sym in "ACGT":
            leaf = {"type": "NucleotideEquals", "position": pos + 1, "symbol": sym}
            c = count(leaf)
            if 0 < c < args.rows // 1000:
                leaves.append((leaf, c))
                break
        if len(leaves) == 4:
            break
    out = {"rows": args.rows, "partitions": len(db.owned), "container_keys": n_keys, "cases": []}
    cases = [("one private mutation", leaves[0][0], [leaves[0]]),
             ("And of two private mutations", {"type": "And", "children": [leaves[0][0], leaves[1][0]]}, leaves[:2]),
             ("Or of four private mutations", {"type": "Or", "children": [l for l, _ in leaves]}, leaves),
             ("True", {"type": "True"}, [])]
    for label, flt, used in cases:
        n, ms = timed(flt)
        # what the reference reads: every leaf's containers (array containers: 2 bytes per value + 8 per container; at most one container per key)
        b_filter = sum(2 * c + 8 * min(c, n_keys) for _, c in used)
        out["cases"].append({"filter": label, "leaf_cardinalities": [c for _, c in used], "count": n, "gpu_ms": ms, "reference_bytes": b_filter,
                             "dense_result_bytes": args.rows // 8, "note": "gpu_ms = CUDA-event time of the prepared query (filter evaluation + rank tables + cardinality), one graph launch"})
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
