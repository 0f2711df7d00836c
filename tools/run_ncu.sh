#!/bin/bash
# one ncu --set full capture of the container-scan kernel (stream and item variants); reports under gpurun_out/
mkdir -p gpurun_out
for mode in stream item; do
  env SB_K1=$mode timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k1_ --launch-skip 30 --launch-count 1 \
      -f -o gpurun_out/k1_$mode python tools/k1_experiment.py --name ncu_$mode --steps 5 > gpurun_out/ncu_$mode.log 2>&1
  tail -2 gpurun_out/ncu_$mode.log
done
