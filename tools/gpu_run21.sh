#!/bin/bash
# measurement: k_viswalk<1024> for few views of a large map (RC_VW_WIDE 0 / 1) on c3, after the tests in $2
T=${1:-r04v}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "$2" > gpurun_out/${T}_pytest_k.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest_k.log
for v in 0 1; do
  RC_VW_WIDE=$v timeout 400 python bench.py --config c3 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_bench_c3_w$v.json 2> gpurun_out/${T}_bench_c3_w$v.err
done
tail -n 3 gpurun_out/${T}_pytest_k.log
