#!/bin/bash
# what the driver runs at round end: smoke, then both bench arms with their defaults
T=${1:-r03b}
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "exit $?" >> gpurun_out/${T}_smoke.log
SECONDS=0; timeout 600 python bench.py --impl reference > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err
echo "ref arm ${SECONDS}s"; SECONDS=0; timeout 600 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
tail -n 3 gpurun_out/${T}_smoke.log; tail -n 2 gpurun_out/${T}_bench_ref.err; tail -n 2 gpurun_out/${T}_bench.err
echo "cuda arm ${SECONDS}s"
