#!/bin/bash
# GPU check of the entity stages: studio / sprite parity tests (per-kernel sync mode first), then the c4 bench line
T=${1:-r02h}
mkdir -p gpurun_out
RC_SYNC_EACH=1 timeout 400 python -m pytest tests -m gpu -q -x -k "studio or sprites or models" > gpurun_out/${T}_pytest_ent.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest_ent.log
timeout 300 python bench.py --config c4 --steps 20 --warmup 3 --cpu-seconds 5 > gpurun_out/${T}_bench_c4.json 2> gpurun_out/${T}_bench_c4.err
timeout 300 python bench.py --config c5 --steps 10 --warmup 3 --cpu-seconds 5 > gpurun_out/${T}_bench_c5.json 2> gpurun_out/${T}_bench_c5.err
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
tail -n 5 gpurun_out/${T}_pytest_ent.log; tail -n 3 gpurun_out/${T}_pytest.log; tail -c 600 gpurun_out/${T}_bench_c4.err
