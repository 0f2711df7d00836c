#!/bin/bash
# GPU round: C4 (AminoAcidMutations over all genes, 10 M rows) with containers up to 16 / 64 / 256 values as postings, then the parity
# suite and the bench line with the default
mkdir -p gpurun_out
TAG=${1:-r02p}
for fm in 16 64 256; do
  SB_FLAT_MAX=$fm timeout -s KILL 900 python tools/configs_report.py --only C4 > gpurun_out/${TAG}_c4_fm$fm.jsonl 2> gpurun_out/${TAG}_c4_fm$fm.err
  python - <<EOF
import json
for l in open("gpurun_out/${TAG}_c4_fm$fm.jsonl"):
    d = json.loads(l); print("flat_max $fm", d["config"], "e2e", round(d["e2e_ms_median"], 3), "device", round(d.get("device_ms_per_query"), 3), "rows", d["result_rows"])
EOF
done
timeout -s KILL 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout -s KILL 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python - <<EOF
import json
d = json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d["roofline"]["avg_launch_ms"], d["query_gpu_ms"], d["throughput"]["value"], d["pool"])
EOF
tail -3 gpurun_out/${TAG}_bench.err
