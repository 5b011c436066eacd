#!/bin/bash
# GPU check: selected tests (-k $2), then bench lines for the configs in $3
T=${1:-r02t}
mkdir -p gpurun_out
RC_SYNC_EACH=0 timeout 600 python -m pytest tests -m gpu -q -x -k "$2" > gpurun_out/${T}_pytest_k.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest_k.log
for c in $3; do
  timeout 400 python bench.py --config $c --steps 30 --warmup 5 --cpu-seconds 5 > gpurun_out/${T}_bench_$c.json 2> gpurun_out/${T}_bench_$c.err
done
tail -n 5 gpurun_out/${T}_pytest_k.log; for c in $3; do tail -c 300 gpurun_out/${T}_bench_$c.err; done
