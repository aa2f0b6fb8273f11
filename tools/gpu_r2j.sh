#!/bin/bash
# round 2, call J (re-entry): where HEAD stands — GPU tests without the long statistics, default bench line, generic quick timings
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q -k "not statistics" > gpurun_out/r2j_pytest.log 2>&1 ) 2>&1 | grep real; echo "pytest exit $?"
tail -30 gpurun_out/r2j_pytest.log
( time timeout 1200 python bench.py > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err ) 2>&1 | grep real; tail -5 gpurun_out/r2j_bench.err
cut -c1-6000 gpurun_out/r2j_bench.json
timeout 900 python tools/quick_gen.py kmer4_200k convex_200k kmer3_wca_500k kmer3_lj_500k > gpurun_out/r2j_quick.log 2>&1; cat gpurun_out/r2j_quick.log
