#!/bin/bash
# round-2 GPU check 1: parity suite + bench lines for c2 / c3 / c5
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02a_smi.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02a_pytest.log
python bench.py --config c2 --steps 50 --warmup 5 > gpurun_out/r02a_bench_c2.json 2> gpurun_out/r02a_bench_c2.err
python bench.py --config c3 --steps 10 --warmup 3 --cpu-seconds 5 > gpurun_out/r02a_bench_c3.json 2> gpurun_out/r02a_bench_c3.err
python bench.py --config c5 --steps 10 --warmup 3 --cpu-seconds 5 > gpurun_out/r02a_bench_c5.json 2> gpurun_out/r02a_bench_c5.err
tail -3 gpurun_out/r02a_pytest.log
