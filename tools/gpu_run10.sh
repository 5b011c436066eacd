#!/bin/bash
# GPU check: the no-alloc test alone in a fresh process, then the whole parity suite (it runs last there)
T=${1:-r02r}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/${T}_pytest_alone.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest_alone.log
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
tail -n 6 gpurun_out/${T}_pytest_alone.log; tail -n 6 gpurun_out/${T}_pytest.log
