#!/bin/bash
# ncu --set full captures of the other kernels of a query (one launch each, steady state); reports under gpurun_out/
mkdir -p gpurun_out
for k in k2_missing interpret_plans; do
  timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 30 --launch-count 1 \
      -f -o gpurun_out/$k python tools/k1_experiment.py --name ncu_$k --steps 5 > gpurun_out/ncu_$k.log 2>&1
  tail -1 gpurun_out/ncu_$k.log | cut -c1-200
done
