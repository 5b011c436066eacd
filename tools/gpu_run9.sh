#!/bin/bash
# GPU check: entity parity tests (per-kernel sync), c4 bench line, c4 launch list
T=${1:-r02m}
mkdir -p gpurun_out
RC_SYNC_EACH=1 timeout 400 python -m pytest tests -m gpu -q -x -k "studio or sprites or models" > gpurun_out/${T}_pytest_ent.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest_ent.log
timeout 300 python bench.py --config c4 --steps 20 --warmup 3 --cpu-seconds 5 > gpurun_out/${T}_bench_c4.json 2> gpurun_out/${T}_bench_c4.err
bash tools/gpu_run7.sh $T c4
tail -n 5 gpurun_out/${T}_pytest_ent.log; tail -c 300 gpurun_out/${T}_bench_c4.err
