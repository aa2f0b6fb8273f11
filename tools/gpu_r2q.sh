#!/bin/bash
# round 2, call Q: cell-list g(r), lattice tests, generic sweep cell-population scan, LJ profile
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_observables.py tests/test_gpu_lattice.py tests/test_gpu_facade.py -q 2>&1 | tail -6
timeout 300 python tools/time_gr.py 2>&1 | tail -5
for tp in 6 10 16 24; do echo "== target population $tp"; RP_SWEEP_TARGET_POP=$tp timeout 600 python tools/quick_gen.py kmer4_200k convex_200k kmer3_wca_500k 2>&1 | tail -3; done
THERM=2 REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^k_gen_sweep" -s 17 -c 1 -o gpurun_out/r2q_prof_gen_lj -f python tools/quick_gen.py kmer3_lj_500k > gpurun_out/r2q_ncu_gen_lj.log 2>&1
python tools/ncu_summary.py gpurun_out/r2q_prof_gen_lj.ncu-rep 30 > gpurun_out/r2q_ncu_gen_lj.txt 2>&1
python tools/ncu_lines.py gpurun_out/r2q_prof_gen_lj.ncu-rep 0 45 >> gpurun_out/r2q_ncu_gen_lj.txt 2>&1
rm -f gpurun_out/r2q_prof_gen_lj.ncu-rep
