#!/bin/bash
# measurement: k_viswalk's shared-memory staging choices (RC_VIS_STAGE bit mask: 1 tree, 2 scratch, 4 lists) on the configs in $3
T=${1:-r03t}
mkdir -p gpurun_out
for m in $2; do
  for c in $3; do
    RC_VIS_STAGE=$m timeout 400 python bench.py --config $c --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_bench_${c}_m$m.json 2> gpurun_out/${T}_bench_${c}_m$m.err
  done
done
ls gpurun_out | grep ${T} | wc -l
