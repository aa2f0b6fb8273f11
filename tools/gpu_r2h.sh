#!/bin/bash
# round 2, call H: parity of the restructured generic sweep + timings
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -k "not statistics" > gpurun_out/r2h_pytest.log 2>&1; echo "pytest exit $?"
tail -30 gpurun_out/r2h_pytest.log
timeout 600 python tools/quick_gen.py kmer4_200k convex_200k kmer3_wca_500k kmer3_lj_500k > gpurun_out/r2h_quick.log 2>&1; cat gpurun_out/r2h_quick.log
COUNTING=1 timeout 600 python tools/quick_gen.py kmer4_200k convex_200k >> gpurun_out/r2h_quick.log 2>&1; tail -2 gpurun_out/r2h_quick.log
