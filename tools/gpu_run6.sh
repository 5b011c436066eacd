#!/bin/bash
# GPU check: parity suite, then the c4 and c2 bench lines
T=${1:-r02j}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
timeout 300 python bench.py --config c4 --steps 20 --warmup 3 --cpu-seconds 5 > gpurun_out/${T}_bench_c4.json 2> gpurun_out/${T}_bench_c4.err
timeout 300 python bench.py --config c2 --steps 50 --warmup 5 --cpu-seconds 5 > gpurun_out/${T}_bench_c2.json 2> gpurun_out/${T}_bench_c2.err
tail -n 12 gpurun_out/${T}_pytest.log; tail -c 400 gpurun_out/${T}_bench_c4.err
