#!/bin/bash
# GPU parity tests only (optionally a -k filter as $1); statistics output kept
mkdir -p gpurun_out
if [ -n "$1" ]; then
  timeout 2800 python -m pytest tests -m gpu -q -s -k "$1" > gpurun_out/pytest_gpu.log 2>&1
else
  timeout 2800 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1
fi
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -n "^\[\|  density\|  g_shell\|  p2\|  tau\|passed\|failed\|exit\|^FAILED\|^E  " gpurun_out/pytest_gpu.log | tail -70
