#!/bin/bash
# full ncu capture of one kernel (regex $3) of one config's resident re-runs; plain run first
T=${1:-r02l}; CFG=${2:-c4}; K=${3:-k_tris_fast}
mkdir -p gpurun_out
timeout 200 python tools/prof.py $CFG > gpurun_out/${T}_prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 1 -c 2 -f -o gpurun_out/${T}_prof python tools/prof.py $CFG > gpurun_out/${T}_ncu.log 2>&1
tail -n 3 gpurun_out/${T}_ncu.log
