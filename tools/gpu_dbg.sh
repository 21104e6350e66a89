#!/bin/bash
export CUDA_LAUNCH_BLOCKING=1 ORBX_GRAPHS=0
echo "== gen1"; ORBX_FAST_GEN=1 python tools/one_frame.py 2>&1 | tail -10
echo "== gen2 no tma"; ORBX_FAST_TMA=0 python tools/one_frame.py 2>&1 | tail -10
echo "== gen2 tma"; python tools/one_frame.py 2>&1 | tail -10
