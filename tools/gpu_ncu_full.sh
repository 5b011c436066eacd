#!/bin/bash
# one full capture of a kernel ($2, regex) of the bench command of config $3 (default: the default bench)
T=${1:-r03g}; K=${2:-k_draw}
C=${3:+--config $3}
mkdir -p gpurun_out
timeout 600 python bench.py $C --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_plain.json 2> gpurun_out/${T}_plain.err && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 8 -c 2 -f -o gpurun_out/${T}_full python bench.py $C --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu.log 2>&1
tail -n 2 gpurun_out/${T}_ncu.log | cut -c1-200
