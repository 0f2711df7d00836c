#!/usr/bin/env python
"""Kernel experiment harness (diagnostics, not part of the product path): times the container scan of the bench query
on a variant build of the libraries (SILO_B200_LIBDIR) and, for a -DSB_K1_TIMELINE build, dumps the per-CTA time stamps.

    SILO_B200_LIBDIR=build_variants/tl python tools/k1_experiment.py --name tl --rows 10000000
"""
import argparse
import ctypes
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "lapis-silo_b200"))
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--name", default="base")
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--out", default=str(ROOT / "gpurun_out"))
    args = ap.parse_args()
    import numpy as np
    import torch
    import pysilo
    import bench
    db = pysilo.Database.synthetic(args.rows, 0x51100003)
    engine = pysilo.Engine(db, 0, db.owned)
    query = bench.c3_query(db)
    table_len = 16 * 29903
    counts = torch.zeros(table_len, dtype=torch.int32, device="cuda:0")
    prepared = pysilo.Prepared(engine, query)
    info = prepared.run_device(0, counts.data_ptr(), table_len, want_bytes=True)
    checksum = int(counts.to(torch.int64).sum().item())
    out = {"name": args.name, "rows": args.rows, "k1_bytes": info["k1_bytes"], "checksum": checksum,
           "env": {k: v for k, v in os.environ.items() if k.startswith("SB_")}}
    for overlap in (False, True):
        pysilo.set_overlap(engine, overlap)
        for _ in range(3):
            prepared.run_device(0, counts.data_ptr(), table_len)
        pysilo.k1_accumulated(engine, reset=True)
        pysilo.timer_begin(engine)
        for _ in range(args.steps):
            prepared.run_device(0, counts.data_ptr(), table_len)
        step_ms = pysilo.timer_end(engine) / args.steps
        k1_ms, n = pysilo.k1_accumulated(engine, reset=True)
        key = "overlap" if overlap else "alone"
        out[f"k1_ms_{key}"] = k1_ms / max(n, 1)
        out[f"step_ms_{key}"] = step_ms
    out["k1_gbs_alone"] = info["k1_bytes"] / (out["k1_ms_alone"] * 1e-3) / 1e9
    out["items"] = pysilo.pool_stats(engine)["n_work_items"]
    lib = pysilo._sb()
    if hasattr(lib, "sb_debug_timeline_enable"):
        pysilo.set_overlap(engine, False)
        blocks = 16384
        assert lib.sb_debug_timeline_enable(ctypes.c_uint32(blocks)) == 0
        prepared.run_device(0, counts.data_ptr(), table_len)
        buf = np.zeros(blocks * 8, dtype=np.uint64)
        lib.sb_debug_timeline_read.argtypes = [ctypes.c_void_p, ctypes.c_uint32]
        assert lib.sb_debug_timeline_read(buf.ctypes.data, blocks) == 0
        buf = buf.reshape(blocks, 8)
        buf = buf[buf[:, 1] != 0]
        Path(args.out).mkdir(exist_ok=True)
        np.save(Path(args.out) / f"timeline_{args.name}.npy", buf)
        out["timeline_blocks"] = int(len(buf))
    print("K1EXP " + json.dumps(out), flush=True)
    prepared.close()


if __name__ == "__main__":
    main()
