#!/bin/bash
T=${1:-r02g}
mkdir -p gpurun_out
RC_SYNC_EACH=1 timeout 300 python -m pytest tests -m gpu -q -x -k "sprites" > gpurun_out/${T}_sprites.log 2>&1; echo "exit $?" >> gpurun_out/${T}_sprites.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
tail -n 5 gpurun_out/${T}_sprites.log gpurun_out/${T}_pytest.log
