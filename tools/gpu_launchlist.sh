#!/bin/bash
# ncu launch list of the bench command (thermalisation shortened); sliced to the warm-up + timed cycles by the caller
mkdir -p gpurun_out
E2E_PIPELINE=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_all.csv python bench.py --configs none --no-cpu-baseline --therm 300 --steps 3 --warmup 3 > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"; wc -l gpurun_out/launches_all.csv
