#!/bin/bash
# the allocation-free k-NN entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
echo "== alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -6
echo "== whole suite"; timeout 2400 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -8 | tee gpurun_out/r2c_pytest_gpu.log
