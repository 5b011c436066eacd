#!/bin/bash
# robustness: the parity suite with a synchronise + error check after every kernel, and again without CUDA graphs
T=${1:-r03x}
mkdir -p gpurun_out
RC_SYNC_EACH=1 timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest_sync.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest_sync.log
RC_NO_GRAPHS=1 timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest_nographs.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest_nographs.log
tail -n 3 gpurun_out/${T}_pytest_sync.log; tail -n 3 gpurun_out/${T}_pytest_nographs.log
