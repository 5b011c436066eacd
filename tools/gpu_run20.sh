#!/bin/bash
# measurement: k_events block size (RC_EV_THREADS) on the configs in $3, after the tests in $4
T=${1:-r04r}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "$4" > gpurun_out/${T}_pytest_k.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest_k.log
for n in $2; do
  for c in $3; do
    RC_EV_THREADS=$n timeout 400 python bench.py --config $c --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_bench_${c}_t$n.json 2> gpurun_out/${T}_bench_${c}_t$n.err
  done
done
tail -n 3 gpurun_out/${T}_pytest_k.log
