#!/bin/bash
# GPU round: timings of the bench queries, ncu launch list, ncu --set full of every kernel of one query (reports under gpurun_out/)
mkdir -p gpurun_out
TAG=${1:-r02}
ROWS=${2:-40000000}
timeout -s KILL 900 python tools/profile_query.py --rows $ROWS --queries 0,1,4 --runs 5 > gpurun_out/${TAG}_times_overlap.json 2> gpurun_out/${TAG}_times.err; tail -c 600 gpurun_out/${TAG}_times_overlap.json
timeout -s KILL 900 python tools/profile_query.py --rows $ROWS --queries 0,1,4 --runs 5 --serial --one-shot > gpurun_out/${TAG}_times_serial.json 2>> gpurun_out/${TAG}_times.err; tail -c 600 gpurun_out/${TAG}_times_serial.json
timeout -s KILL 1200 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches.csv \
    python tools/profile_query.py --rows $ROWS --queries 0 --runs 1 > gpurun_out/${TAG}_ncu_launches.log 2>&1; tail -2 gpurun_out/${TAG}_ncu_launches.log | cut -c1-300
timeout -s KILL 1500 ncu --profile-from-start off --set full --clock-control none --import-source on -f -o gpurun_out/${TAG}_query0 \
    python tools/profile_query.py --rows $ROWS --queries 0 --runs 1 --serial > gpurun_out/${TAG}_ncu_full.log 2>&1; tail -2 gpurun_out/${TAG}_ncu_full.log | cut -c1-300
ls -la gpurun_out
