#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_matcher_gpu.py -x -q -m gpu --timeout 180 > gpurun_out/pytest_matcher.log 2>&1; echo "exit $?" >> gpurun_out/pytest_matcher.log
tail -40 gpurun_out/pytest_matcher.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
