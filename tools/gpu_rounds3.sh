#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_matcher_gpu.py tests/test_cpp_adapter_gpu.py tests/test_agents_gpu.py tests/test_env_switches_gpu.py::test_matcher_resolver_switches_do_not_change_results tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/rounds_tests.log 2>&1; echo "tests rc $?"; tail -5 gpurun_out/rounds_tests.log
python - <<'PY'
import json, sys
sys.argv=['bench.py']
import bench
rows = bench.search_latency_rows(0)
for r in rows['rows']: print(f"{r['member']:48s} gpu {r['gpu_call_ms']:.3f} ms  cpu {r['cpu_port_call_ms']:.3f} ms  x{r['ratio']:.1f}")
PY
