#!/bin/bash
# quick iteration call: GPU tests + dev probe
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
THERM=${THERM:-10} timeout 600 python tools/quick_bench.py 100k 1M 2>&1 | tee gpurun_out/quick.log
