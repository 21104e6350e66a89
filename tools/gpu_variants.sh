#!/bin/bash
for v in build_variants/*.so; do
  cp $v mam3slam_b200/libmam3slam_b200.so
  echo "== $v"; timeout 120 python -m pytest tests/test_matcher_gpu.py -q -m gpu --timeout 60 -k "${KEXPR:-bow}" 2>&1 | tail -3
done
