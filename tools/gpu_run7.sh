#!/bin/bash
# per-kernel launch list (ncu, gpu__time_duration) of one config's resident re-runs; plain run first
T=${1:-r02k}; CFG=${2:-c4}
mkdir -p gpurun_out
timeout 200 python tools/prof.py $CFG > gpurun_out/${T}_prof_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_${CFG}.csv python tools/prof.py $CFG > gpurun_out/${T}_ncu.log 2>&1
tail -n 3 gpurun_out/${T}_prof_plain.log; tail -n 3 gpurun_out/${T}_ncu.log
