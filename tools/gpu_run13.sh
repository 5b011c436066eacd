#!/bin/bash
# ncu launch list of the forced-rebuild C2 batch (cache kernels)
T=${1:-r03a}
mkdir -p gpurun_out
timeout 200 python tools/prof.py c2 rebuild > gpurun_out/${T}_prof_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:k_cache -c 60 --csv --log-file gpurun_out/${T}_cache_launches.csv python tools/prof.py c2 rebuild > gpurun_out/${T}_ncu.log 2>&1
tail -n 2 gpurun_out/${T}_prof_plain.log | cut -c1-300
