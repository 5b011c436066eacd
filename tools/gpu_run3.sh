#!/bin/bash
T=${1:-r02e}
mkdir -p gpurun_out
bash tools/gpu_run1.sh $T
timeout 200 python tools/prof.py c2 rebuild > gpurun_out/${T}_prof_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_draw|k_cache_build' -s 2 -c 4 -f -o gpurun_out/${T}_prof python tools/prof.py c2 rebuild > gpurun_out/${T}_ncu.log 2>&1
tail -n 3 gpurun_out/${T}_ncu.log
