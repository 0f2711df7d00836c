#!/bin/bash
# runs tools/k1_experiment.py over the variant matrix on the GPU box; results in gpurun_out/k1exp.log
mkdir -p gpurun_out
log=gpurun_out/k1exp.log
: > $log
run() { name=$1; shift; echo "== $name" >> $log; env "$@" timeout -s KILL 400 python tools/k1_experiment.py --name $name >> $log 2>&1; }
run stream A=1
run nofence SILO_B200_LIBDIR=build_variants/nofence
run cost12 SILO_B200_LIBDIR=build_variants/cost12
run stream_tl SILO_B200_LIBDIR=build_variants/tl
grep K1EXP $log
