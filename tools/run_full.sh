#!/bin/bash
# full GPU round: parity suite, bench line, ncu launch list + one full capture of the scan kernel, 40M-row line
mkdir -p gpurun_out
timeout -s KILL 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout -s KILL 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 300 gpurun_out/bench.json
timeout -s KILL 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -c 300 gpurun_out/bench_reference.json
timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --in-flight 0 > gpurun_out/ncu_launches.log 2>&1
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k1_ --launch-skip 30 --launch-count 1 -f -o gpurun_out/k1_item python tools/k1_experiment.py --name ncu_item --steps 5 > gpurun_out/ncu_item.log 2>&1
timeout -s KILL 900 python bench.py --rows 40000000 --steps 10 --no-cpu-baseline > gpurun_out/bench_40m.json 2> gpurun_out/bench_40m.err; tail -c 300 gpurun_out/bench_40m.json
