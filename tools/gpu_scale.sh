#!/bin/bash
# the driver's scaling launch at N GPUs (N = $1): torchrun, defaults
N=${1:-2}; T=${2:-r03c}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/${T}_scale_n$N.json 2> gpurun_out/${T}_scale_n$N.err
tail -c 400 gpurun_out/${T}_scale_n$N.err; nvidia-smi topo -m > gpurun_out/${T}_topo_n$N.txt 2>&1
