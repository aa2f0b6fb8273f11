#!/bin/bash
# round 2, call B: whole GPU suite + ncu capture of the restructured k_sweep
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest exit $?"
tail -25 gpurun_out/r2b_pytest.log
CMD="python tools/quick_bench.py 1M"
THERM=10 timeout 300 $CMD > gpurun_out/r2b_plain.log 2>&1 && \
THERM=10 timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_sweep$" -s 40 -c 2 -o gpurun_out/r2b_prof_sweep -f $CMD > gpurun_out/r2b_ncu_full.log 2>&1
echo "ncu full exit $?"
THERM=4 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2b_launches.csv $CMD > gpurun_out/r2b_ncu_list.log 2>&1
echo "ncu list exit $?"
ls -la gpurun_out | tail -5
