#!/bin/bash
# end of round: smoke(), the whole parity suite, the default bench line and the reference arm, as the driver runs them
T=${1:-r04m}
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/${T}_smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_pytest.log
timeout 600 python bench.py --impl reference > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
timeout 600 python bench.py > gpurun_out/${T}_bench_default.json 2> gpurun_out/${T}_bench_default.err
tail -n 2 gpurun_out/${T}_smoke.log; tail -n 3 gpurun_out/${T}_pytest.log; tail -c 300 gpurun_out/${T}_bench_reference.json
