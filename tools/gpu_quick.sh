#!/bin/bash
# quick GPU check: extractor + matcher parity, bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench exit $?"
cat gpurun_out/bench_quick.json; tail -3 gpurun_out/bench_quick.err
