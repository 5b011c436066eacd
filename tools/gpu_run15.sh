#!/bin/bash
# GPU check: selected tests (-k $2), then the c5 / c4 bench lines with k_tris_ordered's strip count varied ($3)
T=${1:-r03i}
mkdir -p gpurun_out
RC_SYNC_EACH=0 timeout 600 python -m pytest tests -m gpu -q -x -k "$2" > gpurun_out/${T}_pytest_k.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest_k.log
for n in $3; do
  for c in c5 c4; do
    RC_TR_STRIPS=$n timeout 400 python bench.py --config $c --steps 20 --warmup 5 --cpu-seconds 2 > gpurun_out/${T}_bench_${c}_s$n.json 2> gpurun_out/${T}_bench_${c}_s$n.err
  done
done
tail -n 5 gpurun_out/${T}_pytest_k.log
