#!/bin/bash
# round 2, call I: parity after the generic-sweep restructuring, timings, 512-thread k_sweep variant on small systems
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -k "not statistics" > gpurun_out/r2i_pytest.log 2>&1; echo "pytest exit $?"
tail -30 gpurun_out/r2i_pytest.log
timeout 900 python tools/quick_gen.py kmer4_200k convex_200k kmer3_wca_500k kmer3_lj_500k > gpurun_out/r2i_quick.log 2>&1; cat gpurun_out/r2i_quick.log
for v in "256 3" "512 1" "512 2" "128 5"; do
  set -- $v
  echo "=== variant threads=$1 minb=$2" >> gpurun_out/r2i_quick_sc.log
  RP_SWEEP_THREADS=$1 RP_SWEEP_MINB=$2 THERM=10 timeout 300 python tools/quick_bench.py 12k 100k >> gpurun_out/r2i_quick_sc.log 2>&1
done
grep -v "count_total\|try_scaling" gpurun_out/r2i_quick_sc.log
