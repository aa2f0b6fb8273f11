#!/bin/bash
# round 2, call E: statistics tests + bench line + ncu captures of the current k_sweep / k_sweep_prepare
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -k "statistics" > gpurun_out/r2e_pytest.log 2>&1; echo "pytest exit $?"
tail -5 gpurun_out/r2e_pytest.log
timeout 600 python bench.py > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err; echo "bench exit $?"; cat gpurun_out/r2e_bench.json
CMD="python tools/quick_bench.py 1M"
THERM=10 timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_sweep$" -s 40 -c 1 -o gpurun_out/r2e_prof_sweep -f $CMD > gpurun_out/r2e_ncu_full.log 2>&1
echo "ncu sweep exit $?"
THERM=10 timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_sweep_prepare" -s 6 -c 1 -o gpurun_out/r2e_prof_prepare -f $CMD > gpurun_out/r2e_ncu_full2.log 2>&1
echo "ncu prepare exit $?"
