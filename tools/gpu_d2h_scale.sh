#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/r03d_d2h_scale.jsonl
for N in 1 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 tools/d2h_scale_probe.py 2>/dev/null | grep n_gpus >> gpurun_out/r03d_d2h_scale.jsonl
done
cat gpurun_out/r03d_d2h_scale.jsonl; nproc; grep MemTotal /proc/meminfo
