#!/bin/bash
# round 2, call P: sanity after the revert, ncu --set full of k_sweep on the thermalised bench state, nematic snapshot at p = 1.4
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -k "not statistics" 2>&1 | tail -6
B="python bench.py --configs none --no-cpu-baseline --therm 300 --steps 3 --warmup 3"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_sweep$" -s 2410 -c 1 -o gpurun_out/r2p_prof_k_sweep -f $B > gpurun_out/r2p_ncu_full.log 2>&1
echo "ncu full exit $?"
python tools/ncu_summary.py gpurun_out/r2p_prof_k_sweep.ncu-rep 40 > gpurun_out/r2p_ncu_k_sweep.txt 2>&1
python tools/ncu_lines.py gpurun_out/r2p_prof_k_sweep.ncu-rep 0 60 >> gpurun_out/r2p_ncu_k_sweep.txt 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_recount_cull$" -s 12 -c 1 -o gpurun_out/r2p_prof_recount -f python tools/quick_bench.py 1M > gpurun_out/r2p_ncu_full2.log 2>&1
python tools/ncu_summary.py gpurun_out/r2p_prof_recount.ncu-rep 30 > gpurun_out/r2p_ncu_recount.txt 2>&1
python tools/ncu_lines.py gpurun_out/r2p_prof_recount.ncu-rep 0 40 >> gpurun_out/r2p_ncu_recount.txt 2>&1
rm -f gpurun_out/r2p_prof_recount.ncu-rep
EQ_TAG=_n14 EQ_POINTS=sc_l5_nematic EQ_PRESSURE_sc_l5_nematic=1.4 EQ_CYCLES=600000 timeout 900 python tests/golden/make_ordered_snapshots.py > gpurun_out/r2p_ordered14.log 2>&1; tail -20 gpurun_out/r2p_ordered14.log | cut -c1-160
