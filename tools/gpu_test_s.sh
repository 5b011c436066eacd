#!/bin/bash
# GPU check of selected tests with output not captured (-s), per-kernel sync mode
T=${1:-r02x}
mkdir -p gpurun_out
RC_SYNC_EACH=1 timeout 600 python -m pytest tests -m gpu -q -x -s -k "$2" > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
grep -n "xh_host\|ref_cuda\|passed\|failed" gpurun_out/${T}_pytest.log | head -20
