#!/bin/bash
# round 2, call C: whole GPU suite (with failure messages) + ncu capture of k_sweep_prepare
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest exit $?"
tail -40 gpurun_out/r2c_pytest.log
CMD="python tools/quick_bench.py 1M"
THERM=10 timeout 300 $CMD > gpurun_out/r2c_plain.log 2>&1 && \
THERM=10 timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_sweep_prepare" -s 6 -c 1 -o gpurun_out/r2c_prof_prepare -f $CMD > gpurun_out/r2c_ncu_full.log 2>&1
echo "ncu full exit $?"
cat gpurun_out/r2c_plain.log
