#!/bin/bash
cp mam3slam_b200/libmam3slam_b200.so /tmp/base.so
for rep in 1 2; do
for lib in /tmp/base.so build_variants/before_block.so; do
  cp $lib mam3slam_b200/libmam3slam_b200.so
  echo "== $lib"; python tools/py_latency.py 2>&1 | tail -3
done; done
cp /tmp/base.so mam3slam_b200/libmam3slam_b200.so
