#!/bin/bash
# the launch list of the bench command of config $2 (default: the default bench), after the same command exited 0 without ncu
T=${1:-r03f}
C=${2:+--config $2}
mkdir -p gpurun_out
timeout 600 python bench.py $C --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_plain.json 2> gpurun_out/${T}_plain.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches.csv python bench.py $C --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu.log 2>&1
tail -c 200 gpurun_out/${T}_plain.err; tail -n 2 gpurun_out/${T}_ncu.log | cut -c1-200; wc -l gpurun_out/${T}_launches.csv
