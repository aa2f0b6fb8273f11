#!/bin/bash
# round 2, call T: the new GPU tests (1e7 snapshot pairs, overlap relaxation, cell-list g(r), smectic statistics with 8 seeds)
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_pairs_snapshot.py tests/test_gpu_relaxation.py tests/test_gpu_observables.py -q -s 2>&1 | tail -12 ) 2>&1 | tail -16
( time timeout 1500 python -m pytest tests/test_gpu_statistics.py -q -s -k "sc_l5_smectic or hs_degenerate" > gpurun_out/r2t_stat.log 2>&1 ) 2>&1 | grep real
grep -v "^$" gpurun_out/r2t_stat.log | tail -14
timeout 300 python tools/time_gr.py > gpurun_out/r2t_time_observables.txt 2>&1; cat gpurun_out/r2t_time_observables.txt
