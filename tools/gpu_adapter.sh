#!/bin/bash
timeout 900 python -m pytest tests/test_cpp_adapter_gpu.py tests/test_zz_no_alloc.py tests/test_matcher_gpu.py -x -q -m gpu -s 2>&1 | grep -v "^$" | tail -8
