#!/bin/bash
# GPU parity tests only (optionally a -k filter as $1)
mkdir -p gpurun_out
if [ -n "$1" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q -k "$1" > gpurun_out/pytest_gpu.log 2>&1
else
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
fi
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log
