#!/bin/bash
# GPU check: the parity suite only (optionally -k EXPR as $2)
T=${1:-r02x}
mkdir -p gpurun_out
if [ -n "$2" ]; then
  RC_SYNC_EACH=1 timeout 600 python -m pytest tests -m gpu -q -x -k "$2" > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
else
  timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
fi
tail -n 30 gpurun_out/${T}_pytest.log
