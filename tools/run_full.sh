#!/bin/bash
# GPU round: parity suite, smoke, bench line
mkdir -p gpurun_out
timeout -s KILL 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout -s KILL 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout -s KILL 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 300 gpurun_out/bench.json
