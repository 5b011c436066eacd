#!/bin/bash
mkdir -p gpurun_out
export RC_SYNC_EACH=1
timeout 120 python tools_gpu_dbg.py tiny 4 > gpurun_out/r02c_tiny.log 2>&1; echo "exit $?" >> gpurun_out/r02c_tiny.log
timeout 120 python tools_gpu_dbg.py small 64 > gpurun_out/r02c_small.log 2>&1; echo "exit $?" >> gpurun_out/r02c_small.log
timeout 150 python tools_gpu_dbg.py large 2 > gpurun_out/r02c_large.log 2>&1; echo "exit $?" >> gpurun_out/r02c_large.log
tail -3 gpurun_out/r02c_tiny.log gpurun_out/r02c_small.log gpurun_out/r02c_large.log
