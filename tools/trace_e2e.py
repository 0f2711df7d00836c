#!/usr/bin/env python
"""SB_TRACE=1 lap times (host engine + library) of end-to-end queries on a synthetic database (diagnostics)."""
import json, os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "lapis-silo_b200")); sys.path.insert(0, str(ROOT))
os.environ["SB_TRACE"] = "1"
import pysilo, bench
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
no_cache = "--no-cache" in sys.argv      # the bench's e2e leg: every call parses, compiles, lowers and submits
db = pysilo.Database.synthetic(rows, 0x51100004, with_amino_acids=True)
engine = pysilo.Engine(db, 0)
if no_cache:
    engine.set_plan_cache(False)
queries = {"C3": bench.c3_query(db),
           "C4": {"action": {"type": "AminoAcidMutations", "minProportion": 0.05}, "filterExpression": {"type": "HasAminoAcidMutation", "sequenceName": "S", "position": 501}}}
for name, q in queries.items():
    text = json.dumps(q).encode()
    for k in range(8 if no_cache else 4):
        sys.stderr.write(f"== {name} call {k}\n"); sys.stderr.flush()
        engine.execute_text(text)
engine.close()
