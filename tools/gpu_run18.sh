#!/bin/bash
# GPU check: the whole parity suite, a subset again with k_viswalk's light staging forced, then bench lines for the configs in $2
T=${1:-r03u}
mkdir -p gpurun_out
S=$SECONDS
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $? after $((SECONDS-S)) s" >> gpurun_out/${T}_pytest.log
RC_VIS_STAGE=2 timeout 900 python -m pytest tests -m gpu -q -x -k "brush or c2 or golden or novis or models" > gpurun_out/${T}_pytest_stage2.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest_stage2.log
for c in $2; do
  timeout 400 python bench.py --config $c --steps 30 --warmup 5 --cpu-seconds 3 > gpurun_out/${T}_bench_$c.json 2> gpurun_out/${T}_bench_$c.err
done
tail -n 4 gpurun_out/${T}_pytest.log; tail -n 3 gpurun_out/${T}_pytest_stage2.log
