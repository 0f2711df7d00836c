#!/bin/bash
# runs tools/k1_experiment.py over the variant matrix on the GPU box; results in gpurun_out/k1exp.log
mkdir -p gpurun_out
log=gpurun_out/k1exp.log
: > $log
run() { name=$1; shift; echo "== $name" >> $log; env "$@" timeout 400 python tools/k1_experiment.py --name $name >> $log 2>&1; }
run base A=1
run pad24 SB_K1_PAD_KB=24
run pad60 SB_K1_PAD_KB=60
run noalloc SILO_B200_LIBDIR=build_variants/noalloc
run noalloc_pad24 SILO_B200_LIBDIR=build_variants/noalloc SB_K1_PAD_KB=24
grep K1EXP $log
