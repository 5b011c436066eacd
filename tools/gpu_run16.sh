#!/bin/bash
# GPU check: the whole parity suite, then bench lines for the configs in $2
T=${1:-r03l}
mkdir -p gpurun_out
S=$SECONDS
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $? after $((SECONDS-S)) s" >> gpurun_out/${T}_pytest.log
for c in $2; do
  timeout 400 python bench.py --config $c --steps 30 --warmup 5 --cpu-seconds 3 > gpurun_out/${T}_bench_$c.json 2> gpurun_out/${T}_bench_$c.err
done
tail -n 6 gpurun_out/${T}_pytest.log
