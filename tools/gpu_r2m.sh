#!/bin/bash
# round 2, call M: lattice population tests, the new statistical cases, nematic starting snapshot at a higher pressure
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lattice.py -q -x 2>&1 | tail -15
( time timeout 2400 python -m pytest tests/test_gpu_statistics.py -q -s -k "not sc_l5_nematic and not sc_l5_smectic" > gpurun_out/r2m_stat.log 2>&1 ) 2>&1 | grep real
grep -v "^$" gpurun_out/r2m_stat.log | tail -60
EQ_TAG=_n16 EQ_POINTS=sc_l5_nematic EQ_CYCLES=400000 timeout 900 python tests/golden/make_ordered_snapshots.py > gpurun_out/r2m_ordered16.log 2>&1; tail -20 gpurun_out/r2m_ordered16.log | cut -c1-160
EQ_TAG=_n19 EQ_POINTS=sc_l5_nematic EQ_PRESSURE_sc_l5_nematic=1.9 EQ_CYCLES=400000 timeout 900 python tests/golden/make_ordered_snapshots.py > gpurun_out/r2m_ordered19.log 2>&1; tail -20 gpurun_out/r2m_ordered19.log | cut -c1-160
