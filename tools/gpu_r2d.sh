#!/bin/bash
# round 2, call D: parity after the k_sweep / k_sweep_prepare changes + timings + launch list
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -k "not statistics" > gpurun_out/r2d_pytest.log 2>&1; echo "pytest exit $?"
tail -30 gpurun_out/r2d_pytest.log
for v in "128 5" "128 6" "128 4"; do
  set -- $v
  echo "=== variant threads=$1 minb=$2" >> gpurun_out/r2d_quick.log
  RP_SWEEP_THREADS=$1 RP_SWEEP_MINB=$2 THERM=10 timeout 300 python tools/quick_bench.py 1M 100k >> gpurun_out/r2d_quick.log 2>&1
done
grep -v "count_total\|try_scaling" gpurun_out/r2d_quick.log
CMD="python tools/quick_bench.py 1M"
THERM=4 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r2d_launches.csv $CMD > gpurun_out/r2d_ncu_list.log 2>&1
echo "ncu list exit $?"
