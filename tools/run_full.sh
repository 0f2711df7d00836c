#!/bin/bash
# full GPU round: parity suite, bench line, host-side lap times of a few steps
mkdir -p gpurun_out
timeout -s KILL 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout -s KILL 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 600 gpurun_out/bench.json
SB_TRACE=1 timeout -s KILL 300 python tools/k1_experiment.py --name trace --steps 3 2>&1 | grep "\[sb\]" | tail -24 > gpurun_out/trace.log
