#!/bin/bash
# ncu full capture of k_sweep on the 1M probe (plain run first, as the recipe demands)
mkdir -p gpurun_out
CMD="python tools/quick_bench.py 1M"
export THERM=5
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k k_sweep -s 20 -c 2 -o gpurun_out/prof_sweep -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_full.log
