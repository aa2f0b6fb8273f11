#!/bin/bash
# round 2, call G: where does the generic sweep spend its time?
mkdir -p gpurun_out
timeout 600 python tools/quick_gen.py kmer4_200k convex_200k kmer3_wca_500k kmer3_lj_500k > gpurun_out/r2g_quick.log 2>&1; cat gpurun_out/r2g_quick.log
THERM=1 REPS=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r2g_launches_kmer4.csv python tools/quick_gen.py kmer4_200k > gpurun_out/r2g_ncu_list.log 2>&1
echo "ncu list exit $?"
THERM=1 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_gen_sweep" -s 9 -c 1 -o gpurun_out/r2g_prof_gen_kmer4 -f python tools/quick_gen.py kmer4_200k > gpurun_out/r2g_ncu_full.log 2>&1
echo "ncu full exit $?"
THERM=1 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_gen_sweep" -s 9 -c 1 -o gpurun_out/r2g_prof_gen_wca -f python tools/quick_gen.py kmer3_wca_500k > gpurun_out/r2g_ncu_full2.log 2>&1
echo "ncu full2 exit $?"
