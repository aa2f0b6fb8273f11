#!/bin/bash
# round 2, call A: parity of the restructured k_sweep + variant timings
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/r2a_host.txt 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/r2a_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q -k "not statistics" > gpurun_out/r2a_pytest.log 2>&1; echo "pytest exit $?"
tail -15 gpurun_out/r2a_pytest.log
for v in "128 4" "128 5" "128 6" "64 8" "64 10" "256 3"; do
  set -- $v
  echo "=== variant threads=$1 minb=$2" >> gpurun_out/r2a_quick.log
  RP_SWEEP_THREADS=$1 RP_SWEEP_MINB=$2 THERM=10 timeout 300 python tools/quick_bench.py 1M 100k >> gpurun_out/r2a_quick.log 2>&1
done
cat gpurun_out/r2a_quick.log | grep -v "count_total\|try_scaling"
timeout 600 compute-sanitizer --tool memcheck python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_memcheck.log 2>&1; tail -3 gpurun_out/r2a_memcheck.log
timeout 900 compute-sanitizer --tool racecheck python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_racecheck.log 2>&1; tail -3 gpurun_out/r2a_racecheck.log
