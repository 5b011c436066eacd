#!/bin/bash
# GPU check: parity suite + bench lines for c2 / c3 / c5 (every command under its own timeout)
T=${1:-r02d}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
timeout 300 python bench.py --config c2 --steps 50 --warmup 5 > gpurun_out/${T}_bench_c2.json 2> gpurun_out/${T}_bench_c2.err
timeout 300 python bench.py --config c3 --steps 10 --warmup 3 --cpu-seconds 5 > gpurun_out/${T}_bench_c3.json 2> gpurun_out/${T}_bench_c3.err
timeout 300 python bench.py --config c5 --steps 10 --warmup 3 --cpu-seconds 5 > gpurun_out/${T}_bench_c5.json 2> gpurun_out/${T}_bench_c5.err
tail -n 3 gpurun_out/${T}_pytest.log
