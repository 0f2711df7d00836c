#!/bin/bash
# GPU round: container-scan variants side by side (one process per packing parameter), results under gpurun_out/
mkdir -p gpurun_out
TAG=${1:-r02}
ROWS=${2:-40000000}
for tile in 256 512; do
  export SB_TILE_POSITIONS=$tile
  timeout -s KILL 900 python tools/profile_query.py --rows $ROWS --queries 0,4 --runs 7 --variants 0,1,2 > gpurun_out/${TAG}_variants_tile$tile.jsonl 2> gpurun_out/${TAG}_variants_tile$tile.err
  echo "== tile $tile"; cut -c1-330 gpurun_out/${TAG}_variants_tile$tile.jsonl; tail -3 gpurun_out/${TAG}_variants_tile$tile.err
done
